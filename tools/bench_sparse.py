#!/usr/bin/env python
"""BASELINE configs[4] shape: fragmented assembly (100 k contigs), read pairs sharded over N GPUs, SPARSE
link tables merged by the hash-partitioned all-to-all (salsa_b200/dist.py).

    python tools/bench_sparse.py [--pairs P]                                   # one GPU
    python -m torch.distributed.run --nproc-per-node 8 ... tools/bench_sparse.py --pairs 500000000

Per step: reset the open-addressing table, ingest the rank's BED (resident in HBM), compact, partition by
owner = mix64(key) % N, all-to-all, add into the owner's table.  Parity at full size without gathering
anything: sum(count) and sum(mix64(key) * count) (mod 2^64) are additive over ranks and keys, so the merged
result and the generators' expected links must agree on both after an all-reduce.  One JSON line (rank 0).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from salsa_b200 import synth  # noqa: E402
from salsa_b200.dist import EngineOps, exchange_sparse  # noqa: E402
from salsa_b200.engine import Engine  # noqa: E402


def mix64_dev(x):
    def shr(v, k):
        return (v >> k) & ((1 << (64 - k)) - 1)
    x = x ^ shr(x, 33)
    x = x * torch.tensor(-49064778989728563, dtype=torch.int64, device=x.device)
    x = x ^ shr(x, 33)
    x = x * torch.tensor(-4265267296055464877, dtype=torch.int64, device=x.device)
    return x ^ shr(x, 33)


def sums(keys, counts):
    return torch.stack([counts.sum(), (mix64_dev(keys) * counts).sum(), torch.tensor(keys.numel(), device=keys.device)])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--pairs', type=int, default=100_000_000, help='pairs per GPU')
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=2)
    a = ap.parse_args()
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (('RANK', 0), ('WORLD_SIZE', 1), ('LOCAL_RANK', 0)))
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    cfg = dict(synth.CONFIGS['cfg5'])
    asm = synth.make_assembly(seed=20261111, **cfg)
    n = asm.ncontig
    buf = torch.empty(int(a.pairs * 2.02 * 48) + (1 << 20), dtype=torch.uint8, device=dev)
    pos, nlines, exp_k, exp_c = 0, 0, [], []
    for done in range(0, a.pairs, 20_000_000):
        m = min(20_000_000, a.pairs - done)
        rec = synth.gen_records(asm, m, seed=(rank + 1) * 1000 + done // 20_000_000, device=dev, read_id_base=done)
        out = synth.format_bed(rec, out=buf[pos:])
        k, c = synth.expected_links(rec, asm.lengths, n)
        exp_k.append(k); exp_c.append(c)
        pos += out.numel(); nlines += rec['ctg'].numel()
        del rec, out
    bed = buf[:pos]
    want = sums(torch.cat(exp_k), torch.cat(exp_c))          # additive: duplicates across chunks need no merging
    del exp_k, exp_c
    eng = Engine(local)
    one = np.ones(n, np.int64)
    eng.set_contigs(asm.names, asm.lengths, one, one)
    ops = EngineOps(eng)
    cap = 1 << 27
    st = torch.cuda.current_stream()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(a.steps)]
    res = {}

    def step(i=None):
        eng.begin_links('sparse', sparse_capacity=cap)
        if i is not None:
            ev[i][0].record(st)
        eng.ingest_device(bed)
        if i is not None:
            ev[i][1].record(st)
        ok, oc, m = eng._compact()
        if world > 1:
            res['k'], res['c'] = exchange_sparse(ok[:m], oc[:m], ops)
        else:
            res['k'], res['c'] = ok[:m], oc[:m]
        if i is not None:
            ev[i][2].record(st)

    for _ in range(a.warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record(st)
    for i in range(a.steps):
        step(i)
    t1.record(st)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    eng.stats_read()
    total_ms = t0.elapsed_time(t1) / a.steps
    ingest_ms = float(np.mean([e[0].elapsed_time(e[1]) for e in ev]))
    merge_ms = float(np.mean([e[1].elapsed_time(e[2]) for e in ev]))
    got = sums(res['k'], res['c'])
    t = torch.tensor([total_ms, ingest_ms, merge_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(got)
        dist.all_reduce(want)
    if rank == 0:
        parity = bool(got[0] == want[0]) and bool(got[1] == want[1])
        line = json.dumps({
            'workload': 'BASELINE configs[4] shape: 100,000 contigs, %d pairs per GPU x %d GPU(s) (%d BED bytes per GPU), sparse '
                        'link tables (capacity 2^27), hash-partitioned all-to-all merge' % (a.pairs, world, bed.numel()),
            'n_gpus': world, 'ms_per_step': float(t[0]), 'ingest_ms': float(t[1]), 'compact_merge_ms': float(t[2]),
            'pairs_per_s': a.pairs * world / (float(t[0]) * 1e-3), 'ingest_gbs_per_gpu': bed.numel() / (float(t[1]) * 1e-3) / 1e9,
            'distinct_links_total': int(got[2]), 'counted_pairs_total': int(got[0]),
            'parity_full_size': 'sum(count) and sum(mix64(key)*count) equal the generators\' expected links: %s' % parity})
        os.write(real_stdout, (line + '\n').encode())
        if not parity:
            sys.exit(1)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
