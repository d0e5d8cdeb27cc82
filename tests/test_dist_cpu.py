"""CPU tests of the multi-GPU host logic: BED sharding (every adjacent-line pair is
evaluated exactly once across shards) and the world_size-2 gloo run of the sparse
merge plumbing (hash partition -> all-to-all -> local reduce) and the dense merge."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _mix64(x):
    x = int(x)
    x ^= x >> 33
    x = x * 0xff51afd7ed558ccd % 2 ** 64
    x ^= x >> 33
    x = x * 0xc4ceb9fe1a85ec53 % 2 ** 64
    x ^= x >> 33
    return x


def test_torch_mix64_matches_the_c_definition():
    import torch
    from salsa_b200.dist import TorchOps
    ks = [0, 1, 2, 12345678901234, 2 ** 62 + 7, 2 ** 63 - 1, 36_000_000 * 4 + 3]
    got = TorchOps.mix64(torch.tensor(ks, dtype=torch.int64)).numpy().view(np.uint64)
    assert [int(g) for g in got] == [_mix64(k) for k in ks]


def test_shard_ranges_cover_every_pair_once(oracle_lib):
    from salsa_b200 import synth
    from salsa_b200.dist import shard_ranges
    asm, rec, bed = synth.small_case(ncontig=25, genome=500_000, npairs=4000, seed=9, triple=0.1, noise=0.3)
    n = asm.ncontig
    one = np.ones(n, np.int64)

    def links(b):
        r = oracle_lib.make_links(b, asm.names, asm.lengths.astype(float), np.ones(n, np.uint8), one, one)
        return {(int(a), int(c), int(e)): int(k) for a, c, e, k in zip(r['first'], r['second'], r['ends'], r['count'])}

    whole = links(bed)
    for world in (1, 2, 3, 8, 64):
        ranges = shard_ranges(bed, world)
        assert len(ranges) == world and ranges[0][0] == 0 and ranges[-1][1] == len(bed)
        merged = {}
        for lo, hi in ranges:
            assert lo == 0 or bed[lo - 1:lo] == b'\n'           # starts on a line start
            for k, v in links(bed[lo:hi]).items():
                merged[k] = merged.get(k, 0) + v
        assert merged == whole, world
    # degenerate inputs
    assert shard_ranges(b'', 4) == [(0, 0)] * 4
    one_line = b'a 1 2 r/1\n'
    assert sum(hi - lo for lo, hi in shard_ranges(one_line, 4)) >= len(one_line)


def test_shard_ranges_with_an_avoid_list(oracle_lib):
    """Iterations >= 2: a skipped (avoid-listed) pair leaves the reference's prev_line stale (make_links.py:72-74), so
    a shard may only start on a line whose contig is in no avoid pair; with `involved` the cuts walk back to one and
    the shards' links add up to the whole file's."""
    from salsa_b200 import synth
    from salsa_b200.dist import shard_ranges
    asm, rec, bed = synth.small_case(ncontig=12, genome=300_000, npairs=6000, seed=21, triple=0.1, noise=0.5)
    n = asm.ncontig
    one = np.ones(n, np.int64)
    avoid = [(asm.names[i], asm.names[j]) for i, j in ((0, 1), (1, 2), (2, 3), (3, 4), (5, 6), (6, 7), (7, 8), (9, 0))]
    involved = set(x for pr in avoid for x in pr)

    def links(b):
        r = oracle_lib.make_links(b, asm.names, asm.lengths.astype(float), np.ones(n, np.uint8), one, one, avoid=avoid)
        return {(int(a), int(c), int(e)): int(k) for a, c, e, k in zip(r['first'], r['second'], r['ends'], r['count'])}

    whole = links(bed)
    assert whole
    differs = False
    for world in (2, 3, 8):
        ranges = shard_ranges(bed, world, involved=involved)
        assert len(ranges) == world and ranges[0][0] == 0 and ranges[-1][1] == len(bed)
        merged = {}
        for lo, hi in ranges:
            assert lo == 0 or bed[lo - 1:lo] == b'\n'
            assert lo == 0 or bed[lo:lo + 64].split(None, 1)[0] not in involved
            for k, v in links(bed[lo:hi]).items():
                merged[k] = merged.get(k, 0) + v
        assert merged == whole, world
        # the cuts that ignore the avoid list are the wrong ones here (that is what the parameter is for)
        naive = {}
        for lo, hi in shard_ranges(bed, world):
            for k, v in links(bed[lo:hi]).items():
                naive[k] = naive.get(k, 0) + v
        differs = differs or naive != whole
    assert differs


def test_fasta_shards_and_merge(oracle_lib):
    """RE counting shards by contig: byte ranges cut at record headers, each counted on its own (here by the oracle),
    merged with the reference's dict semantics -- equal to counting the whole file, for any number of ranks."""
    from salsa_b200 import synth
    from salsa_b200.dist import fasta_shard_ranges, merge_site_counts
    asm = synth.make_assembly(ncontig=23, genome=400_000, len_lo=2_000, len_hi=60_000, nchrom=3, seed=5)
    fa = synth.gen_fasta(asm)
    fa += b'>' + asm.names[3] + b' again\nGATCGATCNNGANTCGAGTC\n'       # a repeated name: first position, last sequence
    whole = oracle_lib.re_sites(fa, 'GATC,GANTC')
    for world in (1, 2, 3, 7, 64):
        ranges = fasta_shard_ranges(fa, world)
        assert len(ranges) == world and ranges[0][0] == 0 and ranges[-1][1] == len(fa)
        parts = []
        for lo, hi in ranges:
            assert hi == lo or fa[lo:lo + 1] == b'>'
            parts.append(oracle_lib.re_sites(fa[lo:hi], 'GATC,GANTC') if hi > lo else ([], [], [], []))
        names, left, right, seqlen = merge_site_counts(parts)
        assert names == list(whole[0]) and (left == whole[1]).all() and (right == whole[2]).all() and (seqlen == whole[3]).all()
    assert fasta_shard_ranges(b'', 3) == [(0, 0)] * 3


def test_coverage_sharded_by_contig_equals_one_run(tmp_path):
    """Coverage shards by contig: the lengths file is dealt to the ranks, each runs the whole program on its contigs
    (here the compiled reference programs), the merged report equals the one-process report."""
    import subprocess
    import cases
    import oracle
    from salsa_b200.dist import merge_coverage_reports, shard_lengths
    if not oracle.ref_binary('break_contigs_start'):
        import pytest
        pytest.skip('oracle/_ref not built')
    d = str(tmp_path)
    seed, lens, holes, npairs = cases.COV_START_CASES['two_contigs']
    names, L, bed = cases.coverage_case(seed, lens, holes, npairs)
    open(d + '/b.bed', 'wb').write(bed)
    ltxt = cases.lengths_text(names, L)
    open(d + '/l.txt', 'wb').write(ltxt)
    run = lambda lp: subprocess.run([oracle.ref_binary('break_contigs_start'), '-a', d + '/b.bed', '-l', lp, '-s', '100'],       # noqa: E731
                                    capture_output=True, check=True).stdout
    whole = run(d + '/l.txt')
    assert whole.strip()
    for world in (2, 3):
        texts = shard_lengths(ltxt, world)
        assert sorted(b''.join(texts).split(b'\n')) == sorted(ltxt.split(b'\n'))
        parts = []
        for r, t in enumerate(texts):
            open(d + '/l%d.txt' % r, 'wb').write(t)
            parts.append(run(d + '/l%d.txt' % r))
        assert sorted(merge_coverage_reports(parts, False).split(b'\n')) == sorted(whole.split(b'\n'))
    # the joins variant: summary lines add up
    seed, lens, holes, npairs, bp = cases.COV_JOIN_CASES['join_ok_and_bad']
    names, L, bed = cases.coverage_case(seed, lens, holes, npairs, scaffold_style=True)
    open(d + '/b.bed', 'wb').write(bed)
    ltxt = cases.lengths_text(names, L)
    open(d + '/l.txt', 'wb').write(ltxt)
    open(d + '/bp.txt', 'wb').write(cases.breakpoints_text({names[k]: v for k, v in bp.items()}))
    runj = lambda lp: subprocess.run([oracle.ref_binary('break_contigs'), '-a', d + '/b.bed', '-l', lp, '-b', d + '/bp.txt', '-i', '2',     # noqa: E731
                                      '-s', '100'], capture_output=True, check=True).stdout
    whole = runj(d + '/l.txt')
    parts = []
    for r, t in enumerate(shard_lengths(ltxt, 2)):
        open(d + '/l%d.txt' % r, 'wb').write(t)
        parts.append(runj(d + '/l%d.txt' % r))
    merged = merge_coverage_reports(parts, True)
    keep = lambda t: sorted(ln for ln in t.split(b'\n') if ln.strip() and len(ln.split()) >= 2)          # noqa: E731
    assert keep(merged) == keep(whole)


WORKER = r'''
import os, sys
sys.path.insert(0, %r)
import numpy as np, torch, torch.distributed as dist
from salsa_b200.dist import TorchOps, exchange_sparse, merge_dense
dist.init_process_group('gloo')
r, w = dist.get_rank(), dist.get_world_size()
rng = np.random.default_rng(100 + r)
# each rank holds partial counts for overlapping key sets
keys = np.unique(rng.integers(0, 5000, 3000)).astype(np.int64)
counts = rng.integers(1, 50, len(keys)).astype(np.int64)
ok, oc = exchange_sparse(torch.from_numpy(keys), torch.from_numpy(counts), TorchOps())
# every key this rank now owns hashes to it, and global totals are preserved
h = TorchOps.mix64(ok).numpy().view(np.uint64) %% np.uint64(w)
assert (h == r).all()
allk = [None] * w; allc = [None] * w; gk = [None] * w; gc = [None] * w
dist.all_gather_object(allk, keys.tolist()); dist.all_gather_object(allc, counts.tolist())
dist.all_gather_object(gk, ok.tolist()); dist.all_gather_object(gc, oc.tolist())
want = {}
for ks, cs in zip(allk, allc):
    for k, c in zip(ks, cs): want[k] = want.get(k, 0) + c
got = {}
for ks, cs in zip(gk, gc):
    for k, c in zip(ks, cs):
        assert k not in got          # owned by exactly one rank
        got[k] = c
assert got == want
# the one-pass route: whole hash tables in, one equal-split all-to-all of segments, owners add
from salsa_b200.dist import TorchSegmentOps, exchange_segments
cap = 8192
tk = np.full(cap, -1, np.int64); tc = np.zeros(cap, np.int64)
slots = rng.choice(cap, len(keys), replace=False)          # any placement: the walk does not depend on the hash
tk[slots] = keys; tc[slots] = counts
sk, sc = exchange_segments(torch.from_numpy(tk), torch.from_numpy(tc), TorchSegmentOps(), slot_cap=4096)
h = TorchOps.mix64(sk).numpy().view(np.uint64) %% np.uint64(w)
assert (h == r).all()
g2k = [None] * w; g2c = [None] * w
dist.all_gather_object(g2k, sk.tolist()); dist.all_gather_object(g2c, sc.tolist())
got2 = {}
for ks, cs in zip(g2k, g2c):
    for k, c in zip(ks, cs):
        assert k not in got2
        got2[k] = c
assert got2 == want
# RE counting sharded by contig: every rank counts its records (here with the oracle), all ranks end with the whole table
import oracle
from salsa_b200 import synth
from salsa_b200.dist import count_sites_sharded
asm = synth.make_assembly(ncontig=17, genome=300_000, len_lo=2_000, len_hi=50_000, nchrom=3, seed=6)
fa = synth.gen_fasta(asm)
got_re = count_sites_sharded(lambda shard, enz: oracle.re_sites(bytes(shard), enz), fa, 'GATC,GANTC')
want_re = oracle.re_sites(fa, 'GATC,GANTC')
assert got_re[0] == list(want_re[0]) and (got_re[1] == want_re[1]).all() and (got_re[2] == want_re[2]).all()
try:
    count_sites_sharded(lambda shard, enz: oracle.re_sites(bytes(shard), enz), fa, 'EcoRI')      # a bad enzyme: raised on every rank
    raise SystemExit('no error')
except (RuntimeError, oracle.OracleError):
    pass
# dense merge: all-reduce(sum)
t = torch.full((64,), r + 1, dtype=torch.int64)
merge_dense(t)
assert int(t[0]) == sum(range(1, w + 1))
dist.destroy_process_group()
sys.stdout.write('rank %%d ok\n' %% r)     # one write per rank: no interleaving
'''


def test_gloo_world2_sparse_and_dense_merge(tmp_path):
    script = tmp_path / 'worker.py'
    script.write_text(WORKER % ROOT)
    r = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr',
                        '127.0.0.1', '--master-port', '29541', str(script)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert 'rank 0 ok' in r.stdout and 'rank 1 ok' in r.stdout
