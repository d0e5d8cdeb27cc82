"""GPU parity tests of the BED ingest path (through the C ABI) against the oracle."""
import gzip
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


@pytest.fixture(scope='module', params=['auto', 'exact', 'cheap'])
def eng(request):
    """One engine per special-byte test of the link scan (picked by a sample of the text / forced exact / forced to
    the cheap superset test): every case below must give the same result in all three."""
    from salsa_b200.engine import Engine
    e = Engine(0)
    e.scan_mode({'auto': 0, 'exact': 1, 'cheap': 2}[request.param])
    yield e
    e.close()


def oracle_dict(o, bed, names, lengths, has_re=None, avoid=()):
    n = len(names)
    hr = np.ones(n, np.uint8) if has_re is None else has_re
    res = o.make_links(bed, names, np.asarray(lengths, float), hr, np.ones(n, np.int64), np.ones(n, np.int64), avoid)
    return {(names[a], names[b], int(e)): int(c) for a, b, e, c in zip(res['first'], res['second'], res['ends'], res['count'])}, res


def gpu_dict(eng, bed, names, lengths, mode='dense', offset=0, has_re=None, via='device', chunks=None):
    n = len(names)
    eng.set_contigs(names, np.asarray(lengths, np.int64), np.ones(n, np.int64), np.ones(n, np.int64), has_re)
    eng.begin_links(mode, sparse_capacity=1 << 16)
    if via == 'device':
        raw = torch.zeros(offset + len(bed) + 64, dtype=torch.uint8, device=eng.device)
        raw[:offset] = 65      # junk before the caller's pointer: must never be read as data
        raw[offset + len(bed):] = 66
        raw[offset:offset + len(bed)] = torch.frombuffer(bytearray(bed), dtype=torch.uint8).to(eng.device) if len(bed) else raw[0:0]
        if chunks is None:
            eng.ingest_device(raw[offset:offset + len(bed)])
        else:
            for lo, hi in chunks:
                eng.ingest_device(raw[offset + lo:offset + hi], base_offset=lo)
    else:
        buf = np.frombuffer(bed, dtype=np.uint8)
        eng.ingest_host(buf, chunk_bytes=via)
    st = eng.stats_read()
    f, s, e, c = eng.links()
    d = {(eng.names[a], eng.names[b], int(x)): int(y) for a, b, x, y in zip(f, s, e, c)}
    assert sum(d.values()) == st.pairs
    return d, st


def test_golden_case(eng, oracle_lib):
    bed = gzip.open(os.path.join(GOLD, 'links_case.bed.gz'), 'rb').read()
    lens = [ln.split() for ln in gzip.open(os.path.join(GOLD, 'links_case.len.gz'), 'rb').read().splitlines()]
    names = [a[0] for a in lens]
    L = [int(a[1]) for a in lens]
    want, res = oracle_dict(oracle_lib, bed, names, L)
    got, st = gpu_dict(eng, bed, names, L)
    assert got == want
    assert st.pairs == res['npairs_counted']
    assert st.lines == bed.count(b'\n') - 1
    # and the committed twin vectors (counts column)
    import json
    g = json.load(open(os.path.join(GOLD, 'links_twin.json')))
    tw = {}
    for ln in g['links']:
        a, b, sc, cnt = ln.split()
        tw[(a.split(':')[0].encode(), b.split(':')[0].encode(), (a[-1] == 'E') * 2 + (b[-1] == 'E'))] = int(cnt)
    assert got == tw


@pytest.mark.parametrize('offset', [0, 1, 7, 15, 16, 33])
def test_misaligned_pointer(eng, oracle_lib, offset):
    from salsa_b200 import synth
    asm, rec, bed = synth.small_case(ncontig=9, genome=100000, npairs=1500, seed=3)
    want, _ = oracle_dict(oracle_lib, bed, asm.names, asm.lengths)
    got, _ = gpu_dict(eng, bed, asm.names, asm.lengths, offset=offset)
    assert got == want


@pytest.mark.parametrize('mode', ['dense', 'sparse'])
@pytest.mark.parametrize('seed', [1, 2, 3])
def test_random_cases(eng, oracle_lib, seed, mode):
    from salsa_b200 import synth
    asm, rec, bed = synth.small_case(ncontig=7 + 30 * seed, genome=400000 * seed, npairs=4000 * seed, seed=10 + seed,
                                     triple=0.05, noise=0.2)
    want, _ = oracle_dict(oracle_lib, bed, asm.names, asm.lengths)
    got, _ = gpu_dict(eng, bed, asm.names, asm.lengths, mode=mode)
    assert got == want
    keys, counts = synth.expected_links(rec, asm.lengths, asm.ncontig)
    assert sum(want.values()) == int(counts.sum())


@pytest.mark.parametrize('seed', [0, 1, 2, 3])
def test_messy_text(eng, oracle_lib, seed):
    rng = np.random.default_rng(seed)
    import cases
    names = cases.MESSY_NAMES
    L = [int(rng.integers(2000, 6000)) for _ in names]
    bed = cases.messy_bed(rng, names, 6000, L) + (b'\n' if seed % 2 else b'')
    want, _ = oracle_dict(oracle_lib, bed, names, L)
    assert len(want) > 50
    got, _ = gpu_dict(eng, bed, names, L)
    assert got == want


def test_many_short_lines_fallback_path(eng, oracle_lib):
    # > 256 line starts inside one warp's 4 KB region -> the 2 KB sub-range path
    rng = np.random.default_rng(5)
    names = [b'a', b'b', b'c']
    L = [10, 10, 10]
    lines = []
    for i in range(40000):
        lines.append(names[int(rng.integers(0, 3))] + b' %d %d r%d' % (rng.integers(0, 9), rng.integers(0, 9), i // 2))
    bed = b'\n'.join(lines) + b'\n'
    want, _ = oracle_dict(oracle_lib, bed, names, L)
    got, st = gpu_dict(eng, bed, names, L)
    assert got == want and st.lines == 39999
    # same contig on both sides never inspects further fields: 1-field lines are fine
    bed2 = b'a\n' * 5000 + b'a 1 2 r/1\nb 3 4 r/2\n'
    want2, _ = oracle_dict(oracle_lib, bed2, names, L)
    got2, _ = gpu_dict(eng, bed2, names, L)
    assert got2 == want2 == {(b'a', b'b', 0): 1}


def test_chunked_equals_whole(eng, oracle_lib):
    from salsa_b200 import synth
    asm, rec, bed = synth.small_case(ncontig=15, genome=300000, npairs=5000, seed=21, triple=0.05)
    want, _ = oracle_dict(oracle_lib, bed, asm.names, asm.lengths)
    # chunk boundaries: each chunk starts at the first byte of the previous chunk's last line
    nl = [i for i, ch in enumerate(bed) if ch == 10]
    cuts = [nl[len(nl) // 5], nl[len(nl) // 2], nl[len(nl) // 2 + 1], nl[-3]]
    chunks, lo = [], 0
    for c in cuts:
        hi = c + 1                                   # chunk ends after line ending at c
        chunks.append((lo, hi))
        lo = bed.rfind(b'\n', 0, c) + 1              # start of that last line
    chunks.append((lo, len(bed)))
    got, _ = gpu_dict(eng, bed, asm.names, asm.lengths, chunks=chunks)
    assert got == want


@pytest.mark.parametrize('chunk', [4096, 65536, 1 << 20])
def test_host_streaming(eng, oracle_lib, chunk):
    from salsa_b200 import synth
    asm, rec, bed = synth.small_case(ncontig=15, genome=300000, npairs=5000, seed=22)
    want, _ = oracle_dict(oracle_lib, bed, asm.names, asm.lengths)
    got, _ = gpu_dict(eng, bed, asm.names, asm.lengths, via=chunk)
    assert got == want


def test_sparse_table_grows_like_the_reference_dict(eng, oracle_lib):
    """A sparse table that is too small is not the input's fault: ingest_links re-ingests with a larger one
    (make_links.py:140-144: the dict just grows) and the result equals the oracle's."""
    from salsa_b200 import synth
    from salsa_b200.engine import DataError
    asm, rec, bed = synth.small_case(ncontig=120, genome=3_000_000, npairs=60000, seed=21, noise=0.5)
    want, _ = oracle_dict(oracle_lib, bed, asm.names, asm.lengths)
    assert len(want) > 3000
    n = asm.ncontig
    eng.set_contigs(asm.names, np.asarray(asm.lengths, np.int64), np.ones(n, np.int64), np.ones(n, np.int64))
    dev_bed = torch.frombuffer(bytearray(bed), dtype=torch.uint8).to(eng.device)
    eng.begin_links('sparse', sparse_capacity=1024)          # far too small: the kernel reports it ...
    eng.ingest_device(dev_bed)
    with pytest.raises(DataError) as ge:
        eng.stats_read()
    assert ge.value.code == 6
    st = eng.ingest_links(lambda: eng.ingest_device(dev_bed), mode='sparse', sparse_capacity=1024)     # ... and this path recovers
    assert eng.sparse[2] >= 4096
    f, s, e, c = eng.links()
    got = {(eng.names[a], eng.names[b], int(x)): int(y) for a, b, x, y in zip(f, s, e, c)}
    assert got == want and sum(got.values()) == st.pairs


def test_100k_contigs_sparse_against_the_oracle(eng, oracle_lib):
    """BASELINE configs[4]'s contig count (dense table impossible: 'auto' must pick the sparse one), 1.5 M pairs:
    the link list equals the oracle's."""
    from salsa_b200 import synth
    asm = synth.make_assembly(ncontig=100_000, genome=300_000_000, len_lo=1_000, len_hi=20_000, nchrom=23, seed=4)
    rec = synth.gen_records(asm, 1_500_000, seed=8, noise=0.05, triple=0.002)
    bed = bytes(synth.format_bed(rec).numpy())
    n = asm.ncontig
    one = np.ones(n, np.int64)
    want = oracle_lib.make_links(bed, asm.names, asm.lengths.astype(float), np.ones(n, np.uint8), one, one)
    eng.set_contigs(asm.names, asm.lengths.astype(np.int64), one, one)
    dev = torch.frombuffer(bytearray(bed), dtype=torch.uint8).to(eng.device)
    st = eng.ingest_links(lambda: eng.ingest_device(dev))
    assert eng.mode == 'sparse'
    f, s2, e, c = eng.links()
    got = {(eng.names[a], eng.names[b], int(x)): int(y) for a, b, x, y in zip(f, s2, e, c)}
    exp = {(asm.names[a], asm.names[b], int(x)): int(y) for a, b, x, y in zip(want['first'], want['second'], want['ends'], want['count'])}
    assert st.pairs == int(want['npairs_counted']) and len(exp) > 100_000
    assert got == exp


@pytest.mark.parametrize('mode', ['dense', 'sparse'])
def test_file_ingest_whole_and_chunked(eng, oracle_lib, tmp_path, mode):
    """Engine.ingest_file: one piece through the pinned readers (the default for a BED that fits the GPU) and the
    line-aligned chunks a larger file is cut into give the oracle's links."""
    from salsa_b200 import synth
    asm, rec, bed = synth.small_case(ncontig=40, genome=800_000, npairs=30000, seed=12, triple=0.05, noise=0.4)
    want, _ = oracle_dict(oracle_lib, bed, asm.names, asm.lengths)
    p = str(tmp_path / 'reads.bed')
    open(p, 'wb').write(bed)
    n = asm.ncontig
    for whole in (None, False):
        eng.set_contigs(asm.names, np.asarray(asm.lengths, np.int64), np.ones(n, np.int64), np.ones(n, np.int64))
        eng.begin_links(mode, sparse_capacity=1 << 16)
        eng.ingest_file(p, chunk_bytes=1 << 18, whole=whole)
        st = eng.stats_read()
        f, s2, e, c = eng.links()
        got = {(eng.names[a], eng.names[b], int(x)): int(y) for a, b, x, y in zip(f, s2, e, c)}
        assert got == want and sum(got.values()) == st.pairs, whole


def test_timers_through_the_abi(eng):
    from salsa_b200 import synth
    asm, rec, bed = synth.small_case(ncontig=30, genome=600_000, npairs=20000, seed=5)
    n = asm.ncontig
    eng.set_contigs(asm.names, np.asarray(asm.lengths, np.int64), np.ones(n, np.int64), np.ones(n, np.int64))
    dev_bed = torch.frombuffer(bytearray(bed), dtype=torch.uint8).to(eng.device)
    eng.timers(enable=True, reset=True)
    eng.begin_links('dense')
    for _ in range(3):
        eng.ingest_device(dev_bed)
    ms, launches = eng.timers(reset=True)
    assert launches == 3 and 0 < ms < 1000
    eng.timers(enable=False)
    eng.ingest_device(dev_bed)
    assert eng.timers() == (0.0, 0)
    eng.stats_read()


def test_empty_and_tiny(eng, oracle_lib):
    names, L = [b'A', b'B'], [1000, 2000]
    for bed in (b'', b'\n', b'A 1 2 r/1', b'A 1 2 r/1\n', b'A 100 200 r1/1\nB 1900 1990 r1/2\n', b'A 100 200 r1/1\nB 1900 1990 r1/2'):
        got, st = gpu_dict(eng, bed, names, L)
        if bed.count(b'B'):
            assert got == {(b'A', b'B', 1): 1}        # KAT-links: A:B  B:E
        else:
            assert got == {}


def test_midpoint_rule(eng):
    names, L = [b'A', b'B'], [1001, 1001]
    for se, e in ((1001, 0), (1002, 3), (1000, 0)):
        bed = b'A 0 %d r/1\nB 0 %d r/2\n' % (se, se)
        got, _ = gpu_dict(eng, bed, names, L)
        assert got == {(b'A', b'B', e): 1}


def test_data_errors(eng, oracle_lib):
    from salsa_b200.engine import DataError
    names, L = [b'A', b'B'], [10, 10]
    cases = [
        (b'A 1 2 r/1\n\nB 1 2 r/2\n', 1, -1),      # blank line: IndexError
        (b'A 1 2 r/1\nB 1 2\n', 1, -1),            # < 4 fields, contigs differ
        (b'A 1 x r/1\nB 1 2 r/2\n', 2, -2),        # ValueError
        (b'A 1 2 r/1\nZ 1 2 r/2\n', 3, -3),        # KeyError (unknown contig)
    ]
    for bed, code, ocode in cases:
        with pytest.raises(oracle_lib.OracleError) as oe:
            oracle_dict(oracle_lib, bed, names, L)
        assert oe.value.code == ocode
        with pytest.raises(DataError) as ge:
            gpu_dict(eng, bed, names, L)
        assert ge.value.code == code
    # zero-length contig, missing RE counts
    with pytest.raises(DataError) as ge:
        gpu_dict(eng, b'A 1 2 r/1\nB 1 2 r/2\n', names, [0, 10])
    assert ge.value.code == 4
    with pytest.raises(DataError) as ge:
        gpu_dict(eng, b'A 1 2 r/1\nB 1 2 r/2\n', names, [10, 10], has_re=np.array([1, 0]))
    assert ge.value.code == 3
    # errors are lazy like the reference: unknown contigs / bad numbers that are never needed are fine
    got, _ = gpu_dict(eng, b'Z x y r/1\nZ 1 2 q/2\nA 1 2 s/1\nQ 1 2 t/1\n', names, L)
    assert got == {}
    # error offset = byte offset of the offending (current) line
    with pytest.raises(DataError) as ge:
        gpu_dict(eng, b'A 1 2 r/1\nA 1 2 r/2\nA 1 x q/1\nB 1 2 q/2\n', names, L)
    assert ge.value.code == 2 and ge.value.offset == 30


def test_line_too_long(eng):
    from salsa_b200.engine import DataError
    names, L = [b'A', b'B'], [10, 10]
    # long TAIL columns are fine as long as the line stays within the halo
    bed = b'A 1 2 r/1 ' + b'x' * 900 + b'\nB 1 2 r/2\n'
    got, _ = gpu_dict(eng, bed, names, L)
    assert got == {(b'A', b'B', 0): 1}
    # a long line that starts less than SALSA_MAX_LINE before a tile boundary is still found
    pad = b'A 1 2 z/1\n' * 3270          # 32.7 KB: the next line straddles byte 32768
    bed = pad + b'A 1 2 r/1 ' + b'x' * 3000 + b'\nB 1 2 r/2\n'
    got, _ = gpu_dict(eng, bed, names, L)
    assert got == {(b'A', b'B', 0): 1}
    # ... but not one that starts more than SALSA_MAX_LINE before it: library limit, reported
    pad = b'A 1 2 z/1\n' * 3070
    bed = pad + b'A 1 2 r/1 ' + b'x' * 3000 + b'\nB 1 2 r/2\n'
    with pytest.raises(DataError) as ge:
        gpu_dict(eng, bed, names, L)
    assert ge.value.code == 5
    # the first four fields themselves running past the halo
    bed = b'A 1 2 z/1\n' * 3276 + b'A 1 2 ' + b'r' * 1100 + b'/1\nB 1 2 ' + b'r' * 1100 + b'/2\n'
    with pytest.raises(DataError) as ge:
        gpu_dict(eng, bed, names, L)
    assert ge.value.code == 5


def test_device_generated_large(eng):
    """2M pairs generated in HBM; the table must equal the generator's own
    statement of the links (never parsed from text)."""
    from salsa_b200 import synth
    asm = synth.make_assembly(ncontig=300, genome=30_000_000, len_lo=20_000, len_hi=400_000, nchrom=5, seed=77)
    rec = synth.gen_records(asm, 2_000_000, seed=78, device='cuda')
    bed = synth.format_bed(rec)
    keys, counts = synth.expected_links(rec, asm.lengths, asm.ncontig)
    n = asm.ncontig
    eng.set_contigs(asm.names, asm.lengths, np.ones(n, np.int64), np.ones(n, np.int64))
    for mode in ('dense', 'sparse'):
        eng.begin_links(mode, sparse_capacity=1 << 20)
        eng.ingest_device(bed)
        st = eng.stats_read()
        f, s, e, c = eng.links()
        k = ((f * n + s) << 2) | e
        assert (k == keys.cpu().numpy()).all() and (c == counts.cpu().numpy()).all()
        assert st.pairs == int(counts.sum()) and st.lines == rec['ctg'].numel() - 1
