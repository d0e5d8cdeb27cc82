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


WORKER = r'''
import os, sys
sys.path.insert(0, %r)
import numpy as np, torch, torch.distributed as dist
import oracle
from salsa_b200 import synth
from salsa_b200.dist import shard_ranges, merge_dense, exchange_sparse, EngineOps
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
dist.barrier()
dist.destroy_process_group()
sys.stdout.write('rank %%d ok\n' %% rank)     # one write per rank: no interleaving
'''


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs two GPUs')
def test_two_rank_nccl_sharded_ingest_and_merge(tmp_path):
    script = tmp_path / 'worker.py'
    script.write_text(WORKER % ROOT)
    r = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr',
                        '127.0.0.1', '--master-port', '29551', str(script)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert 'rank 0 ok' in r.stdout and 'rank 1 ok' in r.stdout
