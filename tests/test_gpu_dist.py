"""GPU tests of the multi-GPU merge: the partition / sparse-add kernels against their torch
statement, and (when the box has two GPUs) a 2-rank NCCL run of sharded ingest + merge."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_partition_and_sparse_add_kernels():
    from salsa_b200.dist import EngineOps, TorchOps
    from salsa_b200.engine import Engine
    eng = Engine(0)
    rng = np.random.default_rng(3)
    keys = np.unique(rng.integers(0, 2 ** 40, 200000)).astype(np.int64)
    counts = rng.integers(1, 1000, len(keys)).astype(np.int64)
    dk, dc = torch.from_numpy(keys).cuda(), torch.from_numpy(counts).cuda()
    for nparts in (1, 2, 3, 8):
        pk, pc, send = EngineOps(eng).partition(dk, dc, nparts)
        tk, tc, tsend = TorchOps().partition(torch.from_numpy(keys), torch.from_numpy(counts), nparts)
        assert send == tsend
        off = 0
        for p in range(nparts):                     # same multiset per part (order inside a part is free)
            a = sorted(zip(pk[off:off + send[p]].cpu().tolist(), pc[off:off + send[p]].cpu().tolist()))
            b = sorted(zip(tk[off:off + send[p]].tolist(), tc[off:off + send[p]].tolist()))
            assert a == b
            off += send[p]
    dup_k = torch.cat([dk, dk[:50000], dk[:10]])
    dup_c = torch.cat([dc, dc[:50000], dc[:10]])
    rk, rc = EngineOps(eng).reduce(dup_k, dup_c)
    wk, wc = TorchOps().reduce(dup_k.cpu(), dup_c.cpu())
    got = dict(zip(rk.cpu().tolist(), rc.cpu().tolist()))
    assert got == dict(zip(wk.tolist(), wc.tolist()))
    eng.close()


def test_segment_partition_and_add_kernels():
    """salsa_sparse_partition / salsa_sparse_add_segments against their numpy statement (dist.TorchSegmentOps), with
    every "rank" emulated on one GPU: the segments of nparts senders are re-assembled by owner and added."""
    from salsa_b200._lib import check
    from salsa_b200.dist import EngineSegmentOps, TorchSegmentOps, segment_bytes
    from salsa_b200.engine import Engine
    eng = Engine(0)
    L, dev = eng.L, eng.device
    st = torch.cuda.current_stream().cuda_stream
    rng = np.random.default_rng(5)
    cap = 1 << 18
    for nparts in (1, 2, 3, 8):
        tables, want = [], {}
        for r in range(nparts):                       # one hash table per emulated sender, overlapping key sets
            keys = np.unique(rng.integers(0, 2 ** 40, 60000)).astype(np.int64)
            counts = rng.integers(1, 1000, len(keys)).astype(np.int64)
            tk = torch.empty(cap, dtype=torch.int64, device=dev)
            tc = torch.empty(cap, dtype=torch.int64, device=dev)
            check(L.salsa_sparse_reset(tk.data_ptr(), tc.data_ptr(), cap, st))
            full = torch.zeros(1, dtype=torch.int32, device=dev)
            dk, dc = torch.from_numpy(keys).to(dev), torch.from_numpy(counts).to(dev)
            check(L.salsa_sparse_add(dk.data_ptr(), dc.data_ptr(), len(keys), tk.data_ptr(), tc.data_ptr(), cap, full.data_ptr(), st))
            assert int(full.item()) == 0
            tables.append((tk, tc))
            for k, c in zip(keys.tolist(), counts.tolist()):
                want[k] = want.get(k, 0) + c
        slot_cap = 65536
        segb = segment_bytes(slot_cap)
        assert L.salsa_sparse_segment_bytes(slot_cap) == segb
        sends = []
        for tk, tc in tables:
            ops = EngineSegmentOps(eng, 1 << 19)
            send, _ = ops.partition(tk, tc, nparts, slot_cap)
            assert ops.flags() == (False, False)
            ref, ovf = TorchSegmentOps().partition(tk.cpu(), tc.cpu(), nparts, slot_cap)
            assert not ovf
            for p in range(nparts):                   # same multiset per segment (order inside a segment is free)
                a = send[p * segb:(p + 1) * segb].cpu().numpy()
                b = ref[p * segb:(p + 1) * segb].numpy()
                na, nb = int(a[:8].view(np.uint64)[0]), int(b[:8].view(np.uint64)[0])
                assert na == nb
                ka, kb = a[16:16 + 8 * na].view(np.int64), b[16:16 + 8 * nb].view(np.int64)
                ca, cb = a[16 + 8 * slot_cap:][:4 * na].view(np.uint32), b[16 + 8 * slot_cap:][:4 * nb].view(np.uint32)
                assert sorted(zip(ka.tolist(), ca.tolist())) == sorted(zip(kb.tolist(), cb.tolist()))
            sends.append(send.clone())
        got = {}
        for owner in range(nparts):                   # what the all-to-all would deliver to `owner`
            recv = torch.cat([s_[owner * segb:(owner + 1) * segb] for s_ in sends])
            ops = EngineSegmentOps(eng, 1 << 19)
            ok, oc = ops.add(recv, nparts, slot_cap)
            assert ops.flags() == (False, False)
            live = (ok != -1) & (oc != 0)
            for k, c in zip(ok[live].cpu().tolist(), oc[live].cpu().tolist()):
                assert k not in got
                got[k] = c
        assert got == want
    # a segment that is too small is reported, never silently short
    ops = EngineSegmentOps(eng, 1 << 19)
    ops.partition(tables[0][0], tables[0][1], 2, 1000)
    assert ops.flags()[0] is True
    eng.close()


def test_re_counting_sharded_by_contig_equals_whole_file():
    """The multi-GPU RE path with its ranks emulated on one GPU: every shard of the FASTA (cut at record headers) goes
    through the single-GPU kernels on its own, the merged table must equal the whole-file run and the oracle."""
    import oracle
    from salsa_b200 import synth
    from salsa_b200.dist import fasta_shard_ranges, merge_site_counts
    from salsa_b200.engine import Engine
    from salsa_b200.re_sites import count_sites
    eng = Engine(0)
    asm = synth.make_assembly(ncontig=31, genome=900_000, len_lo=2_000, len_hi=90_000, nchrom=3, seed=8)
    fa = synth.gen_fasta(asm)
    buf = np.frombuffer(fa, dtype=np.uint8)
    for enz in ('GATC,GANTC', 'AAGCTT', 'DNASE'):
        whole = count_sites(eng, buf, enz)
        for world in (2, 5):
            parts = [count_sites(eng, buf[lo:hi], enz) if hi > lo else ([], [], [], []) for lo, hi in fasta_shard_ranges(fa, world)]
            names, left, right, seqlen = merge_site_counts(parts)
            assert names == list(whole[0]) and (left == whole[1]).all() and (right == whole[2]).all() and (seqlen == whole[3]).all()
        if enz != 'DNASE':
            o = oracle.re_sites(fa, enz)
            assert list(o[0]) == list(whole[0]) and (o[1] == whole[1]).all() and (o[2] == whole[2]).all()
    eng.close()


def test_coverage_sharded_by_contig_equals_one_run(tmp_path):
    """The multi-GPU coverage path with its ranks emulated on one GPU: the drop-in break_contigs_start / break_contigs run on
    every rank's share of the lengths file; the merged report equals the one-process report."""
    import subprocess
    import cases
    from salsa_b200.dist import merge_coverage_reports, shard_lengths
    binp = os.path.join(ROOT, 'salsa_b200', 'bin')
    d = str(tmp_path)
    seed, lens, holes, npairs = cases.COV_START_CASES['two_contigs']
    names, L, bed = cases.coverage_case(seed, lens, holes, npairs)
    open(d + '/b.bed', 'wb').write(bed)
    ltxt = cases.lengths_text(names, L)
    open(d + '/l.txt', 'wb').write(ltxt)
    run = lambda lp: subprocess.run([binp + '/break_contigs_start', '-a', d + '/b.bed', '-l', lp, '-s', '100'],       # noqa: E731
                                    capture_output=True, check=True).stdout
    whole = run(d + '/l.txt')
    assert whole.strip()
    parts = []
    for r, t in enumerate(shard_lengths(ltxt, 2)):
        open(d + '/l%d.txt' % r, 'wb').write(t)
        parts.append(run(d + '/l%d.txt' % r))
    assert sorted(merge_coverage_reports(parts, False).split(b'\n')) == sorted(whole.split(b'\n'))
    seed, lens, holes, npairs, bp = cases.COV_JOIN_CASES['join_ok_and_bad']
    names, L, bed = cases.coverage_case(seed, lens, holes, npairs, scaffold_style=True)
    open(d + '/b.bed', 'wb').write(bed)
    ltxt = cases.lengths_text(names, L)
    open(d + '/l.txt', 'wb').write(ltxt)
    open(d + '/bp.txt', 'wb').write(cases.breakpoints_text({names[k]: v for k, v in bp.items()}))
    runj = lambda lp: subprocess.run([binp + '/break_contigs', '-a', d + '/b.bed', '-l', lp, '-b', d + '/bp.txt', '-i', '2', '-s', '100'],       # noqa: E731
                                     capture_output=True, check=True).stdout
    whole = runj(d + '/l.txt')
    parts = []
    for r, t in enumerate(shard_lengths(ltxt, 2)):
        open(d + '/l%d.txt' % r, 'wb').write(t)
        parts.append(runj(d + '/l%d.txt' % r))
    assert sorted(merge_coverage_reports(parts, True).split(b'\n')) == sorted(whole.split(b'\n'))


WORKER = r'''
import os, sys
sys.path.insert(0, %r)
sys.path.insert(0, os.path.join(sys.path[0], 'tests'))
import numpy as np, torch, torch.distributed as dist
import oracle
from salsa_b200 import synth
from salsa_b200.dist import shard_ranges, merge_dense, exchange_sparse, EngineOps, EngineSegmentOps, exchange_segments
from salsa_b200.engine import Engine
rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dist.init_process_group('nccl', device_id=torch.device('cuda', local))
asm, rec, bed = synth.small_case(ncontig=80, genome=3_000_000, npairs=40000, seed=12, triple=0.05, noise=0.3)
n = asm.ncontig
one = np.ones(n, np.int64)
res = oracle.make_links(bed, asm.names, asm.lengths.astype(float), np.ones(n, np.uint8), one, one)
want = dict(zip((((res['first'].astype(np.int64) * n + res['second']) << 2) | res['ends']).tolist(), res['count'].tolist()))
lo, hi = shard_ranges(bed, world)[rank]
eng = Engine(local)
eng.set_contigs(asm.names, asm.lengths, one, one)
shard = torch.frombuffer(bytearray(bed[lo:hi]), dtype=torch.uint8).cuda()
# dense: all-reduce
eng.begin_links('dense')
eng.ingest_device(shard, base_offset=lo)
eng.stats_read()
merge_dense(eng.table)
nz = torch.nonzero(eng.table).flatten()
assert dict(zip(nz.cpu().tolist(), eng.table[nz].cpu().tolist())) == want
# sparse: partition + all-to-all + add
eng.begin_links('sparse', sparse_capacity=1 << 18)
eng.ingest_device(shard, base_offset=lo)
eng.stats_read()
k, c, m = eng._compact()
ok, oc = exchange_sparse(k[:m], c[:m], EngineOps(eng))
mine = dict(zip(ok.cpu().tolist(), oc.cpu().tolist()))
parts = [None] * world
dist.all_gather_object(parts, mine)
merged = {}
for p in parts:
    for kk, vv in p.items():
        assert kk not in merged
        merged[kk] = vv
assert merged == want
# sparse, one pass and one all-to-all: whole hash tables in, segments exchanged, owners add
keys, cnts, cap = eng.sparse
sops = EngineSegmentOps(eng, 1 << 18)
ok, oc = exchange_segments(keys, cnts, sops, slot_cap=1 << 16)
assert sops.flags() == (False, False)
live = (ok != -1) & (oc != 0)
mine = dict(zip(ok[live].cpu().tolist(), oc[live].cpu().tolist()))
parts = [None] * world
dist.all_gather_object(parts, mine)
merged = {}
for p in parts:
    for kk, vv in p.items():
        assert kk not in merged
        merged[kk] = vv
assert merged == want
# RE counting sharded by contig over the ranks
from salsa_b200.dist import count_sites_sharded
from salsa_b200.re_sites import count_sites
fa = synth.gen_fasta(asm)
fbuf = np.frombuffer(fa, dtype=np.uint8)
got_re = count_sites_sharded(lambda shard, enz: count_sites(eng, shard, enz), fbuf, 'GATC,GANTC')
want_re = oracle.re_sites(fa, 'GATC,GANTC')
assert got_re[0] == list(want_re[0]) and (got_re[1] == want_re[1]).all() and (got_re[2] == want_re[2]).all()
# coverage sharded by contig over the ranks (the library call the drop-in programs wrap)
import ctypes, tempfile, cases
from salsa_b200.dist import coverage_sharded
from salsa_b200 import _lib
wd = tempfile.mkdtemp()
cseed, clens, choles, cpairs = cases.COV_START_CASES['two_contigs']
cnames, cL, cbed = cases.coverage_case(cseed, clens, choles, cpairs)
open(wd + '/b.bed', 'wb').write(cbed)
open(wd + '/l.txt', 'wb').write(cases.lengths_text(cnames, cL))
def run_cov(lp):
    argv = ['-a', wd + '/b.bed', '-l', lp, '-s', '100']
    arr = (ctypes.c_char_p * len(argv))(*[a.encode() for a in argv])
    outp = wd + '/out.rank%%d' %% rank
    assert _lib.lib().salsa_break_contigs_run(0, len(argv), arr, None, 0, outp.encode()) == 0
    return open(outp, 'rb').read()
rep = coverage_sharded(run_cov, wd + '/l.txt', False, wd)
import json
want_cov = json.load(open(os.path.join(%r, 'tests', 'golden', 'coverage_ref.json')))['start']['two_contigs']
assert sorted(rep.decode().split('\n'))[1:] == sorted(want_cov) or sorted(x for x in rep.decode().split('\n') if x) == sorted(want_cov)
dist.barrier()
dist.destroy_process_group()
sys.stdout.write('rank %%d ok\n' %% rank)     # one write per rank: no interleaving
'''


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs two GPUs')
def test_two_rank_nccl_sharded_ingest_and_merge(tmp_path):
    script = tmp_path / 'worker.py'
    script.write_text(WORKER % (ROOT, ROOT))
    r = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr',
                        '127.0.0.1', '--master-port', '29551', str(script)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert 'rank 0 ok' in r.stdout and 'rank 1 ok' in r.stdout
