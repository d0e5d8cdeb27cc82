"""GPU parity tests of the avoid-list path (links_avoid_iteration_N): the reference
`continue`s on an avoided contig pair WITHOUT advancing prev_line
(make_links.py:72-74), so pairs are formed with the last non-skipped line."""
import gzip
import json
import os
import sys
import tempfile

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, 'tests', 'golden')
BIN = os.path.join(ROOT, 'salsa_b200', 'bin')


@pytest.fixture(scope='module')
def eng():
    sys.path.insert(0, BIN)
    from _common import engine
    return engine()


def oracle_dict(o, bed, names, lengths, avoid):
    n = len(names)
    res = o.make_links(bed, names, np.asarray(lengths, float), np.ones(n, np.uint8), np.ones(n, np.int64), np.ones(n, np.int64), avoid)
    return {(names[a], names[b], int(e)): int(c) for a, b, e, c in zip(res['first'], res['second'], res['ends'], res['count'])}


def gpu_dict(eng, bed, names, lengths, avoid, mode='dense', chunks=None):
    n = len(names)
    eng.set_contigs(names, np.asarray(lengths, np.int64), np.ones(n, np.int64), np.ones(n, np.int64), None, avoid)
    eng.begin_links(mode, sparse_capacity=1 << 16)
    dev = torch.frombuffer(bytearray(bed), dtype=torch.uint8).to(eng.device)
    if chunks is None:
        eng.ingest_device(dev)
    else:
        for lo, hi in chunks:
            eng.ingest_device(dev[lo:hi], base_offset=lo)
    eng.stats_read()
    f, s, e, c = eng.links()
    return {(eng.names[a], eng.names[b], int(x)): int(y) for a, b, x, y in zip(f, s, e, c)}


def test_kat_stale_prev(eng, oracle_lib):
    bed = b'A 1 2 r/1\nB 1 2 r/1\nC 1 2 r/2\n'
    names, L = [b'A', b'B', b'C'], [10, 10, 10]
    assert gpu_dict(eng, bed, names, L, []) == {(b'A', b'B', 0): 1, (b'B', b'C', 0): 1}
    assert gpu_dict(eng, bed, names, L, [(b'B', b'A')]) == {(b'A', b'C', 0): 1}
    # an avoid pair naming a contig that is not in the length table still changes what "previous" means
    bed = b'A 1 2 r/1\nZ 1 2 r/1\nC 1 2 r/2\n'
    for av in ([], [(b'A', b'Z')]):
        if not av:
            from salsa_b200.engine import DataError
            with pytest.raises(DataError):            # A-Z counted -> KeyError in the reference
                gpu_dict(eng, bed, names, L, av)
        else:
            assert gpu_dict(eng, bed, names, L, av) == oracle_dict(oracle_lib, bed, names, L, av) == {(b'A', b'C', 0): 1}


def test_golden_avoid_vectors(eng):
    bed = gzip.open(os.path.join(GOLD, 'links_case.bed.gz'), 'rb').read()
    lens = [ln.split() for ln in gzip.open(os.path.join(GOLD, 'links_case.len.gz'), 'rb').read().splitlines()]
    names, L = [a[0] for a in lens], [int(a[1]) for a in lens]
    g = json.load(open(os.path.join(GOLD, 'links_twin.json')))
    avoid = [(a.encode(), b.encode()) for a, b in g['avoid']]
    got = gpu_dict(eng, bed, names, L, avoid)
    tw = {}
    for ln in g['links_avoid']:
        a, b, sc, cnt = ln.split()
        tw[(a.split(':')[0].encode(), b.split(':')[0].encode(), (a[-1] == 'E') * 2 + (b[-1] == 'E'))] = int(cnt)
    assert got == tw                                # the py3 twin's committed vectors


@pytest.mark.parametrize('mode', ['dense', 'sparse'])
@pytest.mark.parametrize('seed,frac', [(1, 0.1), (2, 0.4), (3, 0.8)])
def test_random_avoid_lists(eng, oracle_lib, seed, frac, mode):
    from salsa_b200 import synth
    asm, rec, bed = synth.small_case(ncontig=60, genome=1_500_000, npairs=12000, seed=40 + seed, triple=0.1, noise=0.5)
    rng = np.random.default_rng(seed)
    inv = [asm.names[i] for i in rng.choice(60, int(60 * frac), replace=False)]
    avoid = []
    for _ in range(len(inv) * 2):
        a, b = rng.choice(len(inv), 2, replace=False)
        avoid.append((inv[a], inv[b]))
    want = oracle_dict(oracle_lib, bed, asm.names, asm.lengths, avoid)
    assert want != oracle_dict(oracle_lib, bed, asm.names, asm.lengths, [])
    assert gpu_dict(eng, bed, asm.names, asm.lengths, avoid, mode=mode) == want


def test_long_chains_inside_a_tile(eng, oracle_lib):
    # only 4 contigs, all pairwise avoided except one pair: long runs of skipped lines, but the BED fits one tile
    rng = np.random.default_rng(5)
    names, L = [b'a', b'b', b'c', b'd', b'e'], [100] * 5
    lines = [names[int(rng.integers(0, 5))] + b' %d %d r%d/1' % (rng.integers(0, 99), rng.integers(0, 99), i // 3) for i in range(600)]
    bed = b'\n'.join(lines) + b'\n'
    avoid = [(b'a', b'b'), (b'a', b'c'), (b'b', b'c'), (b'c', b'd'), (b'a', b'd')]
    assert gpu_dict(eng, bed, names, L, avoid) == oracle_dict(oracle_lib, bed, names, L, avoid)


def test_long_avoid_chains_across_tiles(eng, oracle_lib):
    """Every contig avoid-listed and a multi-tile BED: no line before a tile boundary is certainly kept, so the
    walk-back from each tile runs all the way to the buffer's first line (the reference has no limit on such runs,
    make_links.py:72-74; neither has the scanner any more)."""
    names, L = [b'a', b'b', b'c', b'd'], [100] * 4
    rng = np.random.default_rng(6)
    for avoid in ([(b'a', b'b'), (b'b', b'c'), (b'a', b'c')],               # 'd' is free, but rare below
                  [(b'a', b'b'), (b'b', b'c'), (b'a', b'c'), (b'a', b'd'), (b'b', b'd'), (b'c', b'd')]):
        lines = [names[int(rng.integers(0, 3)) if rng.random() < 0.999 else 3] + b' %d %d r%d/1' % (1 + i % 90, 2 + i % 90, i // 2)
                 for i in range(9000)]
        bed = b'\n'.join(lines) + b'\n'
        assert len(bed) > 3 * 32768
        assert gpu_dict(eng, bed, names, L, avoid) == oracle_dict(oracle_lib, bed, names, L, avoid)


def test_file_ingest_when_every_contig_is_avoid_listed(eng, oracle_lib, tmp_path):
    """Engine.ingest_file (what make_links.py -b does): with every scaffold in some avoid pair no line is certainly kept,
    so the file cannot be cut into chunks -- a BED that fits the GPU goes there in one piece and the result is the
    reference's (the step may be slower on such an input, it may not fail)."""
    names, L = [b'a', b'b', b'c', b'd'], [100] * 4
    avoid = [(b'a', b'b'), (b'b', b'c'), (b'a', b'c'), (b'a', b'd'), (b'b', b'd'), (b'c', b'd')]
    rng = np.random.default_rng(16)
    lines = [names[int(rng.integers(0, 4))] + b' %d %d r%d/%d' % (1 + i % 90, 2 + i % 90, i // 2, 1 + i % 2) for i in range(20000)]
    bed = b'\n'.join(lines) + b'\n'
    p = str(tmp_path / 'all_involved.bed')
    open(p, 'wb').write(bed)
    n = len(names)
    for mode in ('dense', 'sparse'):
        eng.set_contigs(names, np.asarray(L, np.int64), np.ones(n, np.int64), np.ones(n, np.int64), None, avoid)
        eng.begin_links(mode, sparse_capacity=1 << 12)
        eng.ingest_file(p, chunk_bytes=1 << 16)
        eng.stats_read()
        f, s, e, c = eng.links()
        assert {(eng.names[a], eng.names[b], int(x)): int(y) for a, b, x, y in zip(f, s, e, c)} == oracle_dict(oracle_lib, bed, names, L, avoid)


def test_host_streaming_with_avoid_list(eng, oracle_lib):
    """salsa_bed_ingest_dense_host with an avoid set: chunks are cut at lines that can never be skipped."""
    from salsa_b200 import synth
    asm, rec, bed = synth.small_case(ncontig=30, genome=600_000, npairs=12000, seed=78, triple=0.1, noise=0.5)
    avoid = [(asm.names[i], asm.names[j]) for i, j in ((0, 1), (2, 3), (4, 5), (1, 6), (7, 0), (8, 9), (10, 11))]
    want = oracle_dict(oracle_lib, bed, asm.names, asm.lengths, avoid)
    n = asm.ncontig
    for chunk in (4096, 65536):
        eng.set_contigs(asm.names, np.asarray(asm.lengths, np.int64), np.ones(n, np.int64), np.ones(n, np.int64), None, avoid)
        eng.begin_links('dense')
        eng.ingest_host(np.frombuffer(bed, dtype=np.uint8), chunk_bytes=chunk)
        eng.stats_read()
        f, s, e, c = eng.links()
        assert {(eng.names[a], eng.names[b], int(x)): int(y) for a, b, x, y in zip(f, s, e, c)} == want


def test_chunks_start_on_kept_lines(eng, oracle_lib):
    from salsa_b200 import synth
    asm, rec, bed = synth.small_case(ncontig=30, genome=600_000, npairs=8000, seed=77, triple=0.1, noise=0.5)
    avoid = [(asm.names[i], asm.names[j]) for i, j in ((0, 1), (2, 3), (4, 5), (1, 6), (7, 0))]
    involved = set(x for p in avoid for x in p)
    want = oracle_dict(oracle_lib, bed, asm.names, asm.lengths, avoid)
    starts = [0] + [i + 1 for i, ch in enumerate(bed) if ch == 10 and i + 1 < len(bed)]
    cuts = []
    for target in (len(starts) // 4, len(starts) // 2, 3 * len(starts) // 4):
        k = target
        while bed[starts[k]:starts[k] + 40].split()[0] in involved:
            k += 1
        cuts.append(k)
    chunks, lo = [], 0
    for k in cuts:
        chunks.append((lo, starts[k + 1]))           # chunk ends after line k ...
        lo = starts[k]                               # ... and the next one starts ON line k (kept for sure)
    chunks.append((lo, len(bed)))
    assert gpu_dict(eng, bed, asm.names, asm.lengths, avoid, chunks=chunks) == want


def test_make_links_script_iteration_2(eng, oracle_lib):
    import make_links
    from salsa_b200 import synth
    asm, rec, bed = synth.small_case(ncontig=25, genome=500_000, npairs=6000, seed=88, noise=0.4)
    dg, do = tempfile.mkdtemp(), tempfile.mkdtemp()
    for d in (dg, do):
        open(d + '/scaffold_length_iteration_2', 'wb').write(asm.lengths_text())
        open(d + '/re_counts_iteration_2', 'wb').write(b''.join(b'%s\t%d\t%d\n' % (n, 3 + i, 40 - i) for i, n in enumerate(asm.names)))
        open(d + '/links_avoid_iteration_2', 'wb').write(b'ctg000001\tctg000002\nctg000003\tctg000007\nnot_a_contig\tctg000005\n')
        open(d + '/b.bed', 'wb').write(bed)
    assert make_links.main(['-b', dg + '/b.bed', '-d', dg, '-i', '2']) == 0
    oracle_lib.make_links_files(do + '/b.bed', do, 2)
    assert open(dg + '/contig_links_iteration_2', 'rb').read() == open(do + '/contig_links_iteration_2', 'rb').read()
