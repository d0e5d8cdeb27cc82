"""GPU parity tests: str(float) round trip, link scores, best-buddy ratios, and the
drop-in make_links.py / fast_scaled_scores.py scripts against the oracle."""
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, 'salsa_b200', 'bin')


@pytest.fixture(scope='module')
def eng():
    sys.path.insert(0, BIN)
    from _common import engine
    return engine()


def fmt(eng, digits, exp10):
    import ctypes as C
    n = len(digits)
    out = C.create_string_buffer(n * 32 + 32)
    off = np.zeros(n + 1, np.int64)
    d = np.ascontiguousarray(digits, np.int64)
    k = np.ascontiguousarray(exp10, np.int32)
    assert eng.L.salsa_format_g12(d.ctypes.data, k.ctypes.data, n, out, n * 32 + 32, off.ctypes.data) == 0
    raw = out.raw
    return [raw[off[i]:off[i + 1]].decode() for i in range(n)]


def test_g12_matches_python2_str(eng, oracle_lib):
    rng = np.random.default_rng(7)
    xs = list(np.exp(rng.uniform(np.log(1e-17), np.log(2e19), 200000)))
    xs += [10.0 ** k for k in range(-17, 20)] + [10.0 ** k * (1 - 2 ** -53) for k in range(-17, 20)]
    xs += [100000000000.5, 100000000001.5, 999999999999.5, 99999999999.95, 0.5, 0.25, 1 / 3.0, 2 / 3.0, 21659 / 210.0,
           3 / 14.0, 123456789012.0, 1234567890123.0, 0.0001, 0.00001, 9.9999999999995e-5, 1e22, 3e-29, 7e29,
           0.0, -2.5, -1 / 3.0]
    xs += [c / (a * 0.5 + b * 0.5) for a, b, c in rng.integers(1, 5000, (50000, 3))]
    # the edges of the 128-bit fast path (1e-8 <= x < 1e12) and exact ties of its rounding step
    xs += [1e-8, 1e-8 * (1 - 2 ** -53), 1e-8 * (1 + 2 ** -52), 9.99999999999e-9, 1e12, 1e12 * (1 - 2 ** -53), 999999999999.75,
           0.5 + 2.0 ** -40, 1.0 + 2.0 ** -36, 2.0 ** -20, 2.0 ** 39 + 0.5, 2.0 ** 38 + 0.5, 1e-29, 1e29]
    xs += [float(m) / 8.0 + 1e5 for m in range(0, 64)] + [(2 * m + 1) / 2.0 ** 14 for m in range(40)]
    x = np.array(xs, dtype=np.float64)
    d, k, w = eng.g12(x)
    txt = fmt(eng, d, k)
    for i in range(0, len(x), 1):
        want = oracle_lib.py2_str_float(float(x[i])) if i % 17 == 0 or i > 200000 else None
        if want is not None:
            assert txt[i] == want, (x[i], txt[i], want)
    # vectorised check of every element against Python's own formatting
    want_all = ['%.12g' % v for v in x]
    want_all = [s if ('.' in s or 'e' in s or 'n' in s) else s + '.0' for s in want_all]
    assert txt == want_all
    assert (w == np.array([float(s) for s in want_all])).all()


def test_g12_values_without_digits(eng):
    """inf / nan print as Python prints them; a finite value outside [1e-30, 1e30] is refused by the formatters instead
    of being printed wrongly (the pipeline's own scores never get there; an arbitrary contig_links file can)."""
    import ctypes as C
    x = np.array([float('inf'), float('-inf'), float('nan'), 2.5], dtype=np.float64)
    d, k, w = eng.g12(x)
    assert fmt(eng, d, k) == ['inf', '-inf', 'nan', '2.5']
    assert np.isinf(w[0]) and np.isinf(w[1]) and np.isnan(w[2]) and w[3] == 2.5
    for bad in (1e-31, 1e31, 5e-324, -3e200):
        d, k, w = eng.g12(np.array([1.5, bad]))
        assert d[1] == np.iinfo(np.int64).min and w[1] == bad
        out = C.create_string_buffer(128)
        off = np.zeros(3, np.int64)
        rc = eng.L.salsa_format_g12(np.ascontiguousarray(d).ctypes.data, np.ascontiguousarray(k).ctypes.data, 2, out, 128, off.ctypes.data)
        assert rc != 0 and b'1e-30' in eng.L.salsa_last_error()


def _case(seed, ncontig=40, npairs=20000):
    from salsa_b200 import synth
    asm, rec, bed = synth.small_case(ncontig=ncontig, genome=20000 * ncontig, npairs=npairs, seed=seed, triple=0.02, noise=0.15)
    rng = np.random.default_rng(seed)
    rl = rng.integers(0, 60, ncontig)
    rr = rng.integers(0, 60, ncontig)
    rl[:3] = 0
    rr[:3] = 0          # norm == 0 -> the int literal 0
    return asm, bed, rl, rr


def _write_dir(d, asm, bed, rl, rr, it=1):
    open(d + '/scaffold_length_iteration_%d' % it, 'wb').write(asm.lengths_text())
    open(d + '/re_counts_iteration_%d' % it, 'wb').write(b''.join(b'%s %d %d\n' % (n, a, b) for n, a, b in zip(asm.names, rl, rr)))
    open(d + '/b.bed', 'wb').write(bed)


@pytest.mark.parametrize('seed', [1, 2])
def test_dropin_scripts_match_oracle(eng, oracle_lib, seed):
    import make_links
    import fast_scaled_scores
    asm, bed, rl, rr = _case(seed)
    dg, do = tempfile.mkdtemp(), tempfile.mkdtemp()
    for d in (dg, do):
        _write_dir(d, asm, bed, rl, rr)
    assert make_links.main(['-b', dg + '/b.bed', '-d', dg, '-i', '1', '-x', 'abc']) == 0
    oracle_lib.make_links_files(do + '/b.bed', do, 1)
    a = open(dg + '/contig_links_iteration_1', 'rb').read()
    b = open(do + '/contig_links_iteration_1', 'rb').read()
    assert a == b and len(a) > 1000
    assert b'\t0\t' in a                     # a norm == 0 link is present and printed as the int 0
    assert fast_scaled_scores.main(['-d', dg, '-i', '1']) == 0
    oracle_lib.fast_scaled_scores_files(do, 1)
    a = open(dg + '/contig_links_scaled_iteration_1', 'rb').read()
    b = open(do + '/contig_links_scaled_iteration_1', 'rb').read()
    assert a == b
    env = dict(os.environ, LC_ALL='C')
    sa = subprocess.run(['sort', '-k', '5', '-gr', dg + '/contig_links_scaled_iteration_1'], capture_output=True, env=env).stdout
    sb = subprocess.run(['sort', '-k', '5', '-gr', do + '/contig_links_scaled_iteration_1'], capture_output=True, env=env).stdout
    assert sa == sb


def test_scaled_kats(eng, oracle_lib):
    import fast_scaled_scores
    cases = [
        ['a:B\tb:B\t4.0\t7\n', 'a:B\tc:B\t2.0\t3\n'],
        ['a:B\tb:B\t3.5\t7\n', 'c:E\td:B\t0\t1\n'],
        ['a:B\tb:B\t2.0\t1\n', 'a:B\tc:B\t2.0\t1\n'],
        ['a:B\tb:B\t2.0\t1\n', 'a:B\tc:B\t2.0\t1\n', 'a:B\td:E\t2.0\t9\n', 'b:B\td:E\t0.5\t2\n'],
        ['x:E\ty:B\t103.138095238\t21659\n', 'x:E\tz:B\t0.214285714286\t1\n', 'y:B\tz:B\t1e-05\t1\n'],
    ]
    rng = np.random.default_rng(3)
    big = []
    for i in range(300):
        for j in rng.choice(300, 4, replace=False):
            if i < j:
                big.append('n%d:%s\tn%d:%s\t%s\t%d\n' % (i, 'BE'[i % 2], j, 'BE'[j % 2],
                                                     oracle_lib.py2_str_float(float(rng.integers(0, 40)) / float(rng.integers(1, 9))),
                                                     rng.integers(1, 500)))
    cases.append(big)
    for lines in cases:
        want = ''.join(oracle_lib.scaled_lines_from_link_lines(lines)).encode()
        got = fast_scaled_scores.scaled_text(eng, [ln.encode() for ln in lines])
        assert got == want


def test_kat_link_score_line(eng):
    # KAT-links: A:B  B:E  0.285714285714  1
    import make_links
    d = tempfile.mkdtemp()
    open(d + '/scaffold_length_iteration_1', 'w').write('A\t1000\nB\t2000\n')
    open(d + '/re_counts_iteration_1', 'w').write('A 3 5\nB 2 4\n')
    open(d + '/b.bed', 'w').write('A 100 200 r1/1\nB 1900 1990 r1/2\n')
    assert make_links.main(['-b', d + '/b.bed', '-d', d, '-i', '1']) == 0
    assert open(d + '/contig_links_iteration_1').read() == 'A:B\tB:E\t0.285714285714\t1\n'


def test_make_links_fails_like_reference(eng):
    import make_links
    from salsa_b200.engine import DataError
    d = tempfile.mkdtemp()
    open(d + '/scaffold_length_iteration_1', 'w').write('A\t1000\n')
    open(d + '/re_counts_iteration_1', 'w').write('A 3 5\n')
    open(d + '/b.bed', 'w').write('A 100 200 r1/1\nB 1900 1990 r1/2\n')
    with pytest.raises(DataError):
        make_links.main(['-b', d + '/b.bed', '-d', d, '-i', '1'])
    assert not os.path.exists(d + '/contig_links_iteration_1')


def test_dropin_scripts_match_reference_vectors(eng):
    """tests/golden/links_ref.json = output of the reference's OWN make_links.py / fast_scaled_scores.py /
    RE_sites.py (run by tests/golden/ref_py2emu.py in the build container): the drop-in scripts must
    write the same lines -- link order is Python-2 dict order in the reference, so sorted lines."""
    import gzip
    import io
    import json
    import make_links
    import fast_scaled_scores
    import RE_sites
    gold = os.path.join(ROOT, 'tests', 'golden')
    gz = lambda fn: gzip.open(os.path.join(gold, fn), 'rb').read()        # noqa: E731
    g = json.load(open(os.path.join(gold, 'links_ref.json')))
    lines = lambda p: sorted(open(p).read().splitlines(keepends=True))    # noqa: E731
    d = tempfile.mkdtemp()
    open(d + '/a.fa', 'wb').write(gz('links_case.fa.gz'))
    for enz in ('GATC', 'GATC,GANTC', 'DNASE', 'AAGCTT'):
        buf = io.StringIO()
        assert RE_sites.main(['-a', d + '/a.fa', '-e', enz], out=buf) == 0
        assert buf.getvalue().splitlines() == g['re_' + enz]
    open(d + '/scaffold_length_iteration_1', 'wb').write(gz('links_case.len.gz'))
    open(d + '/re_counts_iteration_1', 'w').write('\n'.join(g['re_GATC,GANTC']) + '\n')
    open(d + '/b.bed', 'wb').write(gz('links_case.bed.gz'))
    for f in ('scaffold_length', 're_counts'):
        os.link(d + '/%s_iteration_1' % f, d + '/%s_iteration_2' % f)
    open(d + '/links_avoid_iteration_2', 'w').write(''.join('%s\t%s\n' % tuple(p) for p in g['avoid']))
    for it, lk, sk in ((1, 'links', 'scaled'), (2, 'links_avoid', 'scaled_avoid')):
        assert make_links.main(['-b', d + '/b.bed', '-d', d, '-i', str(it)]) == 0
        assert lines('%s/contig_links_iteration_%d' % (d, it)) == g[lk]
        assert fast_scaled_scores.main(['-d', d, '-i', str(it)]) == 0
        assert lines('%s/contig_links_scaled_iteration_%d' % (d, it)) == g[sk]
    m = tempfile.mkdtemp()
    open(m + '/scaffold_length_iteration_1', 'wb').write(gz('messy_case.len.gz'))
    open(m + '/re_counts_iteration_1', 'wb').write(gz('messy_case.re.gz'))
    open(m + '/b.bed', 'wb').write(gz('messy_case.bed.gz'))
    assert make_links.main(['-b', m + '/b.bed', '-d', m, '-i', '1']) == 0
    assert lines(m + '/contig_links_iteration_1') == g['messy_links']
    assert fast_scaled_scores.main(['-d', m, '-i', '1']) == 0
    assert lines(m + '/contig_links_scaled_iteration_1') == g['messy_scaled']


@pytest.mark.parametrize('seed', [1, 2, 3])
def test_gpu_sort_equals_coreutils_sort(eng, seed):
    """The sorted file the in-process pipeline leaves behind must be what `sort -k 5 -gr` (LC_ALL=C) writes:
    ratio descending as a number, equal ratios by the whole line, bytewise, descending."""
    import fast_scaled_scores
    rng = np.random.default_rng(seed)
    # names with prefix relations (ctg1 / ctg10 / ctg1_2), long names, and many exact ties (best buddies have ratio 1.0)
    names = ['ctg%d' % i for i in range(40)] + ['ctg1_%d' % i for i in range(5)] + ['a_very_long_contig_name_%03d' % i for i in range(15)]
    lines, seen = [], set()
    for _ in range(3000):
        a, b = (int(x) for x in rng.integers(0, len(names), 2))
        if a == b or (min(a, b), max(a, b)) in seen:
            continue
        seen.add((min(a, b), max(a, b)))
        w = [1.0, 0.5, 2.0, 1e-05, 12345.678, float(rng.integers(1, 50)) / float(rng.integers(1, 9))][int(rng.integers(0, 6))]
        lines.append(('%s:%s\t%s:%s\t%s\t%d\n' % (names[a], 'BE'[int(rng.integers(2))], names[b], 'BE'[int(rng.integers(2))],
                                               ('%.12g' % w) if '.' in ('%.12g' % w) or 'e' in ('%.12g' % w) else ('%.12g' % w) + '.0',
                                               int(rng.integers(1, 999)))).encode())
    # an edge may appear once per (u, v) NODE pair; drop duplicates of node pairs
    uniq, out = set(), []
    for ln in lines:
        a = ln.split()
        k = tuple(sorted((a[0], a[1])))
        if k not in uniq and a[0] != a[1]:
            uniq.add(k)
            out.append(ln)
    text, srt = fast_scaled_scores.scaled_text(eng, out, want_sorted=True)
    want = subprocess.run(['sort', '-k', '5', '-gr'], input=text, capture_output=True, env=dict(os.environ, LC_ALL='C')).stdout
    assert srt == want and len(srt) > 10000
    ratios = [ln.split(b'\t')[4] for ln in srt.splitlines()]
    assert max(ratios.count(r) for r in set(ratios)) > 50          # the tie rule really was exercised
