"""CPU tests of the oracle itself: known-answer tests derived by hand from the
reference source, the py3 twin's golden vectors, and the
compiled reference's golden outputs for coverage."""
import gzip
import json
import os
import subprocess
import tempfile

import numpy as np
import pytest

import cases

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def gz(name):
    return gzip.open(os.path.join(GOLD, name), 'rb').read()


def test_py2_str_float(oracle_lib):
    o = oracle_lib
    # the only expected values in the reference tree: fast_scaled_scores.py:39-40
    assert o.py2_str_float(21659 / 210.0) == '103.138095238'
    assert o.py2_str_float(3 / 14.0) == '0.214285714286'
    assert o.py2_str_float(1.0) == '1.0'
    assert o.py2_str_float(0.0) == '0.0'
    assert o.py2_str_float(1e16) == '1e+16'
    assert o.py2_str_float(1e-5) == '1e-05'
    assert o.py2_str_float(123456789012.0) == '123456789012.0'
    assert o.py2_str_float(1234567890123.0) == '1.23456789012e+12'
    assert o.py2_str_float(2.5) == '2.5'
    from oracle import py_twin
    rng = np.random.default_rng(0)
    for x in np.exp(rng.uniform(-20, 20, 500)):
        assert o.py2_str_float(float(x)) == py_twin.py2_str(float(x))
        assert o.round12(float(x)) == float(py_twin.py2_str(float(x)))


def _links(o, bed, names, lengths, rl, rr, avoid=()):
    res = o.make_links(bed, names, np.asarray(lengths, float), np.ones(len(names), np.uint8), rl, rr, avoid)
    return o.link_lines(res, names)


def test_kat_links(oracle_lib):
    # KAT-links: make_links.py:81-144,165-169
    bed = b'A 100 200 r1/1\nB 1900 1990 r1/2\n'
    lines = _links(oracle_lib, bed, [b'A', b'B'], [1000, 2000], [3, 2], [5, 4])
    assert lines == ['A:B\tB:E\t0.285714285714\t1\n']


def test_kat_midpoint(oracle_lib):
    # B iff pos <= len/2 (make_links.py:108), E iff pos > len - len/2 (:122)
    names, L = [b'A', b'B'], [1001, 1001]
    for se, exp in ((1001, 'B'), (1002, 'E'), (1000, 'B')):
        bed = b'A 0 %d r/1\nB 0 %d r/2\n' % (se, se)
        lines = _links(oracle_lib, bed, names, L, [1, 1], [1, 1])
        assert lines[0].startswith('A:%s\tB:%s\t' % (exp, exp))


def test_kat_sliding_window_and_order(oracle_lib):
    # three records of one read: (1,2) and (2,3) are both evaluated (make_links.py:60-146);
    # first/second ordered by byte-string '<' with positions swapped (:90-96)
    bed = b'b 0 10 r/1\nB 90 100 r/1\nb 90 100 r/2\nx 1 2 q/1\n'
    lines = _links(oracle_lib, bed, [b'b', b'B', b'x'], [100, 100, 100], [1, 1, 1], [2, 2, 2])
    # 'B' < 'b': pair1 first=B(pos 95 -> E) second=b(pos 5 -> B); pair2 first=B(E) second=b(pos 95 -> E)
    assert sorted(lines) == ['B:E\tb:B\t0.666666666667\t1\n', 'B:E\tb:E\t0.5\t1\n']


def test_kat_norm_zero_and_errors(oracle_lib):
    o = oracle_lib
    bed = b'A 1 2 r/1\nB 1 2 r/2\n'
    assert _links(o, bed, [b'A', b'B'], [10, 10], [0, 0], [0, 0]) == ['A:B\tB:B\t0\t1\n']   # int 0, make_links.py:165-166
    with pytest.raises(o.OracleError) as e:
        _links(o, bed, [b'A'], [10], [0], [0])
    assert e.value.code == -3       # KeyError
    with pytest.raises(o.OracleError) as e:
        _links(o, b'A 1 2 r/1\n\nB 1 2 r/2\n', [b'A', b'B'], [10, 10], [0, 0], [0, 0])
    assert e.value.code == -1       # IndexError on the blank line
    with pytest.raises(o.OracleError) as e:
        _links(o, b'A 1 x r/1\nB 1 2 r/2\n', [b'A', b'B'], [10, 10], [0, 0], [0, 0])
    assert e.value.code == -2
    with pytest.raises(o.OracleError) as e:
        _links(o, bed, [b'A', b'B'], [0, 10], [0, 0], [0, 0])
    assert e.value.code == -4
    # fields are only inspected when needed: same contig -> nothing parsed
    assert _links(o, b'A x\nA y\n', [b'A'], [10], [0], [0]) == []
    # empty input, single line
    assert _links(o, b'', [b'A'], [10], [0], [0]) == []
    assert _links(o, b'A 1 2 r/1', [b'A'], [10], [0], [0]) == []


def test_kat_avoid_stale_prev(oracle_lib):
    # make_links.py:72-74: an avoided pair `continue`s WITHOUT prev_line = line,
    # so line 3 is compared with line 1
    bed = b'A 1 2 r/1\nB 1 2 r/1\nC 1 2 r/2\n'
    names, L = [b'A', b'B', b'C'], [10, 10, 10]
    plain = _links(oracle_lib, bed, names, L, [1, 1, 1], [1, 1, 1])
    assert sorted(plain) == ['A:B\tB:B\t1.0\t1\n', 'B:B\tC:B\t1.0\t1\n']
    av = _links(oracle_lib, bed, names, L, [1, 1, 1], [1, 1, 1], avoid=[(b'B', b'A')])
    assert av == ['A:B\tC:B\t1.0\t1\n']


def test_kat_re(oracle_lib):
    o = oracle_lib
    # RE_sites.py:66-74
    n, l, r, s = o.re_sites(b'>c1 x\nGATCGATC\n', 'GATC')
    assert (n, list(l), list(r), list(s)) == ([b'c1'], [1], [1], [8])
    # all Ns replaced by the same base (RE_sites.py:52-56): GANNTC -> GAGGTC GAAATC GATTTC GACCTC
    n, l, r, s = o.re_sites(b'>a\nGAGGTCGAAATCGATTTCGACCTCGAGATC\n', 'GANNTC')
    assert int(l[0] + r[0]) == 4
    # non-overlapping finditer: AAAA on AAAAAAAAA -> starts {0,4}
    n, l, r, s = o.re_sites(b'>a\nAAAAA\nAAAA\n', 'AAAA')
    assert (int(l[0]), int(r[0])) == (1, 1)
    # duplicates in the enzyme list are counted twice; lower case / N never match
    n, l, r, s = o.re_sites(b'>a\nGATCgatcGANC\n', 'GATC,GATC')
    assert (int(l[0]), int(r[0])) == (2, 0)
    # DNASE: len/2 both (RE_sites.py:39-43); multi-line, trailing whitespace stripped
    n, l, r, s = o.re_sites(b'>a d\nACGTA \nCC\n>b\nA\n', 'DNASE')
    assert (list(l), list(r)) == ([3, 0], [3, 0])
    with pytest.raises(o.OracleError):
        o.re_sites(b'>a:b\nAC\n', 'GATC')
    with pytest.raises(o.OracleError):
        o.re_sites(b'>a\nAC\n', 'MboI')


def test_kat_scaled(oracle_lib):
    # fast_scaled_scores.py:64-86
    lines = ['a:B\tb:B\t4.0\t7\n', 'a:B\tc:B\t2.0\t3\n']
    out = oracle_lib.scaled_lines_from_link_lines(lines)
    assert out == ['a:B\tb:B\t4.0\t2.0\t2.0\t7\tFalse\tFalse\n', 'a:B\tc:B\t2.0\t4.0\t0.5\t3\tFalse\tFalse\n']
    # lone edge: bestAlt 0 -> int 1 (:83-84); zero weight prints 0.0
    out = oracle_lib.scaled_lines_from_link_lines(['a:B\tb:B\t3.5\t7\n', 'c:E\td:B\t0\t1\n'])
    assert out == ['a:B\tb:B\t3.5\t1\t3.5\t7\tFalse\tFalse\n', 'c:E\td:B\t0.0\t1\t0.0\t1\tFalse\tFalse\n']
    # tie for the best weight at a node: the alternative IS the tied weight
    out = oracle_lib.scaled_lines_from_link_lines(['a:B\tb:B\t2.0\t1\n', 'a:B\tc:B\t2.0\t1\n'])
    assert [ln.split('\t')[3:5] for ln in out] == [['2.0', '1.0'], ['2.0', '1.0']]


def test_kat_sort_ties():
    # `sort -k 5 -gr`: equal ratios ordered by the whole line, descending bytes (run_pipeline.py:188)
    lines = b'a\tb\t1\t1\t1.0\t5\tFalse\tFalse\nc\td\t1\t1\t1.0\t5\tFalse\tFalse\nb\tz\t9\t3\t3.0\t1\tFalse\tFalse\n'
    r = subprocess.run(['sort', '-k', '5', '-gr'], input=lines, capture_output=True, env={'LC_ALL': 'C', 'PATH': os.environ['PATH']})
    assert [x.split(b'\t')[0] for x in r.stdout.splitlines()] == [b'b', b'c', b'a']


def test_golden_twin_vectors(oracle_lib):
    o = oracle_lib
    g = json.load(open(os.path.join(GOLD, 'links_twin.json')))
    bed, fa = gz('links_case.bed.gz'), gz('links_case.fa.gz')
    for enz in ('GATC', 'GATC,GANTC', 'DNASE', 'AAGCTT'):
        n, l, r, s = o.re_sites(fa, enz)
        assert ['%s %d %d' % (a.decode(), b, c) for a, b, c in zip(n, l, r)] == g['re_' + enz]
    d = tempfile.mkdtemp()
    open(d + '/scaffold_length_iteration_1', 'wb').write(gz('links_case.len.gz'))
    open(d + '/re_counts_iteration_1', 'w').write('\n'.join(g['re_GATC,GANTC']) + '\n')
    open(d + '/b.bed', 'wb').write(bed)
    lines = o.make_links_files(d + '/b.bed', d, 1)
    assert sorted(lines) == g['links']
    assert sorted(o.fast_scaled_scores_files(d, 1)) == g['scaled']
    for f in ('scaffold_length', 're_counts'):
        os.link(d + '/%s_iteration_1' % f, d + '/%s_iteration_2' % f)
    open(d + '/links_avoid_iteration_2', 'w').write(''.join('%s\t%s\n' % tuple(p) for p in g['avoid']))
    assert sorted(o.make_links_files(d + '/b.bed', d, 2)) == g['links_avoid']
    assert g['links_avoid'] != g['links']


def test_threads_match_single(oracle_lib):
    from salsa_b200 import synth
    asm, rec, bed = synth.small_case(ncontig=20, genome=300000, npairs=5000, seed=5)
    one = np.ones(20, np.uint8)
    rl = np.arange(20) + 1
    a = oracle_lib.make_links(bed, asm.names, asm.lengths.astype(float), one, rl, rl * 2)
    for nt in (2, 3, 8):
        b = oracle_lib.make_links(bed, asm.names, asm.lengths.astype(float), one, rl, rl * 2, nthreads=nt)
        for k in ('first', 'second', 'ends', 'count'):
            assert (a[k] == b[k]).all()
    keys, counts = synth.expected_links(rec, asm.lengths, 20)
    k2 = ((a['first'].astype(np.int64) * 20 + a['second']) << 2) | a['ends']
    o = np.argsort(k2)
    assert (k2[o] == keys.numpy()).all() and (a['count'][o] == counts.numpy()).all()


def _cov_start_lines(o, names, L, bed):
    return sorted('%s\t%d' % (names[c].decode(), p) for c, p in o.cov_start(bed, names, L))


def _cov_join_lines(o, names, L, bed, bpn):
    rows, tot, sus = o.cov_joins(bed, names, L, bpn)
    out = [names[c].decode() + ''.join('\t%d' % p for p in ps) for c, ps in rows]
    pct = subprocess.run(['awk', 'BEGIN{printf "%g", ' + '%d*100.0/%d' % (sus, tot) + '}'], capture_output=True).stdout.decode() if tot else 'nan'
    out += ['Total Joins = %d' % tot, 'Suspicious Joins = %d' % sus, 'Percent Suspicious Joins = ' + pct]
    return sorted(out)


@pytest.mark.parametrize('case', sorted(cases.COV_START_CASES))
def test_cov_start_golden(oracle_lib, case):
    g = json.load(open(os.path.join(GOLD, 'coverage_ref.json')))
    seed, lens, holes, npairs = cases.COV_START_CASES[case]
    names, L, bed = cases.coverage_case(seed, lens, holes, npairs)
    assert _cov_start_lines(oracle_lib, names, L, bed) == g['start'][case]
    ref = oracle_lib.ref_binary('break_contigs_start')
    if ref:   # live check against the compiled UNMODIFIED reference when it travelled with the repo
        d = tempfile.mkdtemp()
        open(d + '/b.bed', 'wb').write(bed)
        open(d + '/l.txt', 'wb').write(cases.lengths_text(names, L))
        r = subprocess.run([ref, '-a', d + '/b.bed', '-l', d + '/l.txt', '-s', '100'], capture_output=True, check=True)
        assert sorted(r.stdout.decode().splitlines()) == g['start'][case]


@pytest.mark.parametrize('case', sorted(cases.COV_JOIN_CASES))
def test_cov_joins_golden(oracle_lib, case):
    g = json.load(open(os.path.join(GOLD, 'coverage_ref.json')))
    seed, lens, holes, npairs, bp = cases.COV_JOIN_CASES[case]
    names, L, bed = cases.coverage_case(seed, lens, holes, npairs, scaffold_style=True)
    bpn = {names[k]: v for k, v in bp.items()}
    assert _cov_join_lines(oracle_lib, names, L, bed, bpn) == g['joins'][case]


def _ref_dirs(g, kind):
    """-> directory prepared with the committed inputs of a links_ref.json case."""
    d = tempfile.mkdtemp()
    if kind == 'case':
        open(d + '/scaffold_length_iteration_1', 'wb').write(gz('links_case.len.gz'))
        open(d + '/re_counts_iteration_1', 'w').write('\n'.join(g['re_GATC,GANTC']) + '\n')
        open(d + '/b.bed', 'wb').write(gz('links_case.bed.gz'))
        for f in ('scaffold_length', 're_counts'):
            os.link(d + '/%s_iteration_1' % f, d + '/%s_iteration_2' % f)
        open(d + '/links_avoid_iteration_2', 'w').write(''.join('%s\t%s\n' % tuple(p) for p in g['avoid']))
    else:
        open(d + '/scaffold_length_iteration_1', 'wb').write(gz('messy_case.len.gz'))
        open(d + '/re_counts_iteration_1', 'wb').write(gz('messy_case.re.gz'))
        open(d + '/b.bed', 'wb').write(gz('messy_case.bed.gz'))
    return d


def test_golden_reference_vectors(oracle_lib):
    """links_ref.json was produced by the reference's own scripts (tests/golden/ref_py2emu.py): the C
    oracle and the py3 twin must both reproduce it, line for line."""
    from oracle import py_twin
    o = oracle_lib
    g = json.load(open(os.path.join(GOLD, 'links_ref.json')))
    fa = gz('links_case.fa.gz')
    for enz in ('GATC', 'GATC,GANTC', 'DNASE', 'AAGCTT'):
        n, l, r, s = o.re_sites(fa, enz)
        assert ['%s %d %d' % (a.decode(), b, c) for a, b, c in zip(n, l, r)] == g['re_' + enz]
    for impl in ('c', 'twin'):
        d = _ref_dirs(g, 'case')
        if impl == 'c':
            links = lambda it: sorted(o.make_links_files(d + '/b.bed', d, it))        # noqa: E731
            scaled = lambda it: sorted(o.fast_scaled_scores_files(d, it))             # noqa: E731
        else:
            def links(it):
                ls = py_twin.make_links(d + '/b.bed', d, it)
                open('%s/contig_links_iteration_%d' % (d, it), 'w').write(''.join(ls))
                return sorted(ls)
            scaled = lambda it: sorted(py_twin.fast_scaled_scores(d, it))             # noqa: E731
        assert links(1) == g['links']
        assert scaled(1) == g['scaled']
        assert links(2) == g['links_avoid']
        assert scaled(2) == g['scaled_avoid']
        d = _ref_dirs(g, 'messy')
        assert links(1) == g['messy_links']
        assert scaled(1) == g['messy_scaled']
