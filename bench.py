#!/usr/bin/env python
"""bench.py -- Hi-C pairs/s into the contig-end link table (BASELINE.json's metric).

    python bench.py --gpus N --steps K --warmup W           # this repo's CUDA path
    python bench.py --impl reference ...                     # the reference's CPU path

A step = one pass of the hot path (BED text -> link table) over one synthetic
BED of the named shape, already resident in HBM when the timed region starts.
Workload at N=1: BASELINE.json configs[1] (3.1 Gb assembly, 3,000 contigs,
500 M pairs).  For N>1 every rank holds its own 500 M-pair shard of read pairs
(weak scaling) and the per-GPU tables are merged by an NCCL all-reduce inside
the step.  The same JSON line also carries `sparse`: BASELINE configs[4]'s
shape (100,000 contigs, per-GPU SPARSE tables, one hash-partitioned NCCL
all-to-all, local add), timed the same way.  One JSON line is printed by rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

METRIC = 'hic_pairs_per_sec_into_link_table'
UNIT = 'pairs/s'
_REAL_STDOUT = None


def emit(line):
    """The one JSON line goes to the process's real stdout; main() moved fd 1 to stderr so that nothing a
    library prints there (NCCL's version banner) can end up next to it."""
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, (line + '\n').encode())


def ncu_traffic_ratio():
    """dram bytes (read + write) / algorithmic bytes of the ingest kernel, from the committed
    `ncu --set full` capture (profiles/r02_v8_ingest_metrics.txt, 50 M-pair launch = 4,703,357,292 B)."""
    p = os.path.join(ROOT, 'profiles', 'r02_v8_ingest_metrics.txt')
    try:
        rd = wr = None
        for ln in open(p):
            a = ln.split()
            if a and a[0] == 'dram__bytes_read.sum':
                rd = float(a[2]) * {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1.0}[a[1]]
            if a and a[0] == 'dram__bytes_write.sum':
                wr = float(a[2]) * {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1.0}[a[1]]
        return (rd + wr) / 4703357292.0
    except Exception:
        return None


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        return json.load(open(p))['hbm_gbs'], 'measured'
    return 6650.0, 'fallback'


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ('timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '50'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.rows.append((time.time(), ln.strip()))

    def mark(self):
        """start of the timed region (the sampler itself starts earlier: nvidia-smi needs ~0.2 s to come up)"""
        self.t0 = time.time()

    def stop(self):
        if not self.proc:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        t1 = time.time()
        for ts, r in self.rows:
            if ts < getattr(self, 't0', 0) or ts > t1:
                continue
            a = [x.strip() for x in r.split(',')][1:]
            if len(a) < 7:
                continue
            try:
                sm.append(float(a[0])); mx.append(float(a[1]))
            except ValueError:
                continue
            for nm, v in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), a[3:7]):
                if v.lower().startswith('active'):
                    reasons.add(nm)
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': sorted(reasons), 'samples': len(sm)}


def mix64_dev(x):
    """salsa_mix64 (common.cuh) on an int64 tensor (wraps like uint64)."""
    def shr(v, k):
        return (v >> k) & ((1 << (64 - k)) - 1)
    x = x ^ shr(x, 33)
    x = x * torch.tensor(-49064778989728563, dtype=torch.int64, device=x.device)
    x = x ^ shr(x, 33)
    x = x * torch.tensor(-4265267296055464877, dtype=torch.int64, device=x.device)
    return x ^ shr(x, 33)


def build_workload(cfg_name, pairs, device, seed, dense_expected=True):
    """Synthetic assembly + BED text generated straight into HBM, plus the
    generator's own statement of the expected links: the dense table, or (for the
    100,000-contig shape, whose dense table would not fit) two checksums that are
    additive over ranks and keys: sum(count) and sum(mix64(key) * count) mod 2^64."""
    from salsa_b200 import synth
    cfg = dict(synth.CONFIGS[cfg_name])
    asm = synth.make_assembly(seed=20261106 + int(cfg_name[-1]), **cfg)
    n = asm.ncontig
    chunk = 20_000_000
    est = int(pairs * 2.02 * 48) + (1 << 20)
    buf = torch.empty(est, dtype=torch.uint8, device=device)
    expected = torch.zeros(n * n * 4, dtype=torch.int64, device=device) if dense_expected else torch.zeros(2, dtype=torch.int64, device=device)
    pos, done, nlines = 0, 0, 0
    while done < pairs:
        m = min(chunk, pairs - done)
        rec = synth.gen_records(asm, m, seed=seed * 1000 + done // chunk, device=device, read_id_base=done)
        out = synth.format_bed(rec, out=buf[pos:])
        keys, counts = synth.expected_links(rec, asm.lengths, n)
        if dense_expected:
            expected.index_add_(0, keys, counts)
        else:
            expected += torch.stack([counts.sum(), (mix64_dev(keys) * counts).sum()])
        pos += out.numel()
        nlines += rec['ctg'].numel()
        done += m
        del rec, out, keys, counts
    torch.cuda.synchronize()
    return asm, buf[:pos], expected, nlines


def sparse_leg(args, eng, dev, rank, world):
    """BASELINE configs[4]'s shape on this rank's GPU: 100,000 contigs, `--sparse-pairs` pairs of BED text in HBM, an
    open-addressing link table, and the merge north_star names: every link goes to owner = mix64(key) % N in one pass
    over the table, ONE all-to-all with equal splits (NCCL over NVLink), owners add (salsa_b200/dist.exchange_segments).
    Timed like the main step: device events, max over ranks.  Parity at full size without gathering anything: the two
    additive checksums of the merged tables against the generators'."""
    import torch.distributed as dist
    from salsa_b200.dist import EngineSegmentOps, exchange_segments
    pairs = args.sparse_pairs
    asm, bed, want, nlines = build_workload('cfg5', pairs, dev, seed=rank + 1, dense_expected=False)
    n = asm.ncontig
    one = np.ones(n, np.int64)
    eng.set_contigs(asm.names, asm.lengths, one, one)
    cap = 1 << 27
    while cap < pairs // 2:               # counted pairs per rank are ~0.3 per pair on this shape, distinct links fewer: load <= 0.6
        cap <<= 1
    st = torch.cuda.current_stream()
    # size the segments once, from the first ingest (outside the timed region): distinct links / N, plus slack
    eng.begin_links('sparse', sparse_capacity=cap)
    eng.ingest_device(bed)
    eng.stats_read()
    keys, cnts, cap = eng.sparse
    live_t = ((keys != -1) & (cnts != 0)).sum()
    if world > 1:
        dist.all_reduce(live_t, op=dist.ReduceOp.MAX)       # every rank must use the same segment size
    live = int(live_t.item())
    slot_cap = (int(live / world * 1.15) + 4096) // 4 * 4
    cap = outcap = 1 << 20
    while cap < 2 * live:                 # tables at load <= 0.5: this rank's, and its share of the merged one
        cap <<= 1                         # (an owner ends up with about `live` links: weak scaling)
    outcap = cap
    eng.begin_links('sparse', sparse_capacity=cap)
    ops = EngineSegmentOps(eng, outcap)
    nsteps = max(2, min(args.steps, 5))
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(nsteps)]
    res = {}

    def step(i=None):
        keys, cnts, cap = eng.sparse
        eng.L.salsa_sparse_reset(keys.data_ptr(), cnts.data_ptr(), cap, st.cuda_stream)
        if i is not None:
            ev[i][0].record(st)
        eng.ingest_device(bed)
        if i is not None:
            ev[i][1].record(st)
        res['k'], res['c'] = exchange_segments(keys, cnts, ops, slot_cap)
        if i is not None:
            ev[i][2].record(st)

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record(st)
    for i in range(nsteps):
        step(i)
    t1.record(st)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    eng.stats_read()                      # raises on any data error (table full included)
    assert ops.flags() == (False, False), 'sparse merge: a segment or an owner table overflowed'
    t = torch.tensor([t0.elapsed_time(t1) / nsteps, float(np.mean([e[0].elapsed_time(e[1]) for e in ev])),
                      float(np.mean([e[1].elapsed_time(e[2]) for e in ev]))], device=dev, dtype=torch.float64)
    livem = (res['k'] != -1) & (res['c'] != 0)
    got = torch.stack([res['c'][livem].sum(), (mix64_dev(res['k'][livem]) * res['c'][livem]).sum(), livem.sum()])
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(got)
        dist.all_reduce(want)
    parity = bool(got[0] == want[0]) and bool(got[1] == want[1])
    if not parity:
        raise SystemExit('PARITY FAILURE (sparse leg): merged tables differ from the generators\' expected links')
    peak, _ = peaks()
    rec = {'workload': 'BASELINE configs[4] shape: 100,000 contigs, %d pairs per GPU x %d GPU(s) (%d BED bytes per GPU in HBM), sparse '
                       'link tables (2^%d slots), merge = one pass over the table into per-owner segments (12 B per link) + ONE '
                       'equal-split NCCL all-to-all + add into the owner\'s table' % (pairs, world, bed.numel(), cap.bit_length() - 1),
           'ms_per_step': float(t[0]), 'ingest_ms': float(t[1]), 'merge_ms': float(t[2]), 'steps': nsteps,
           'pairs_per_s': pairs * world / (float(t[0]) * 1e-3),
           'frac_of_aggregate_hbm': bed.numel() * world / (float(t[0]) * 1e-3) / 1e9 / (peak * world),
           'ingest_frac_of_hbm': bed.numel() / (float(t[1]) * 1e-3) / 1e9 / peak,
           'distinct_links_total': int(got[2]), 'counted_pairs_total': int(got[0]),
           'exchange_bytes_per_gpu': int(world * (16 + 12 * slot_cap)),
           'parity': parity, 'parity_how': 'sum(count) and sum(mix64(key)*count) mod 2^64 of the merged tables == the generators\' expected links'}
    del bed, ops
    return rec


def small_merge_check(eng, dev, rank, world):
    """A small BED sharded over the ranks, ingested into sparse tables and merged by the same one-pass route; the
    union of what the owners hold must equal the oracle's link table of the whole BED (the oracle only checks)."""
    import torch.distributed as dist
    import oracle
    from salsa_b200 import synth
    from salsa_b200.dist import EngineSegmentOps, exchange_segments, shard_ranges
    asm, rec, bed = synth.small_case(ncontig=80, genome=3_000_000, npairs=40000, seed=12, triple=0.05, noise=0.3)
    n = asm.ncontig
    one = np.ones(n, np.int64)
    eng.set_contigs(asm.names, asm.lengths, one, one)
    lo, hi = shard_ranges(bed, world)[rank]
    eng.begin_links('sparse', sparse_capacity=1 << 18)
    eng.ingest_device(torch.frombuffer(bytearray(bed[lo:hi]), dtype=torch.uint8).to(dev), base_offset=lo)
    eng.stats_read()
    keys, cnts, cap = eng.sparse
    ops = EngineSegmentOps(eng, 1 << 18)
    ok, oc = exchange_segments(keys, cnts, ops, slot_cap=1 << 16)
    assert ops.flags() == (False, False)
    live = (ok != -1) & (oc != 0)
    mine = dict(zip(ok[live].cpu().tolist(), oc[live].cpu().tolist()))
    parts = [None] * world
    if world > 1:
        dist.all_gather_object(parts, mine)
    else:
        parts = [mine]
    if rank != 0:
        return True
    oracle.build()
    res = oracle.make_links(bed, asm.names, asm.lengths.astype(float), np.ones(n, np.uint8), one, one)
    want = dict(zip((((res['first'].astype(np.int64) * n + res['second']) << 2) | res['ends']).tolist(), res['count'].tolist()))
    merged = {}
    for p in parts:
        for k, v in p.items():
            assert k not in merged, 'a key is owned by two ranks'
            merged[k] = v
    if merged != want:
        raise SystemExit('PARITY FAILURE: %d-rank sparse merge differs from the oracle' % world)
    return True


def run_ours(args):
    import torch.distributed as dist
    from salsa_b200.engine import Engine
    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local = int(os.environ.get('LOCAL_RANK', 0))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    pairs = args.pairs
    asm, bed, expected, nlines = build_workload(args.config, pairs, dev, seed=rank + 1)
    n = asm.ncontig
    eng = Engine(local)
    one = np.ones(n, np.int64)
    eng.set_contigs(asm.names, asm.lengths, one, one)
    eng.begin_links('dense')
    st = torch.cuda.current_stream()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]

    def step(i=None):
        eng.table.zero_()
        if i is not None:
            ev[i][0].record(st)
        eng.ingest_device(bed)
        if i is not None:
            ev[i][1].record(st)
        if world > 1:
            dist.all_reduce(eng.table)      # merge per-GPU link tables over NVLink

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.mark()
    torch.cuda.profiler.start()          # lets `ncu --profile-from-start off` see only the timed region
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    t0.record(st)
    for i in range(args.steps):
        step(i)
    t1.record(st)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    total_ms = t0.elapsed_time(t1)
    kern_ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([total_ms, kern_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, kern_ms = float(t[0]), float(t[1])
    # parity at full size: the table must equal the generator's statement of the links
    stt = eng.stats_read()
    if world > 1:
        dist.all_reduce(expected)
    parity = bool(torch.equal(eng.table, expected))
    if not parity and not os.environ.get('SALSA_BENCH_ABLATION'):      # (set only by tools/sweep_stream.sh for kernels with parts compiled out)
        raise SystemExit('PARITY FAILURE: link table differs from the generator\'s expected table')

    # end to end through the C ABI with HOST buffers (pinned), bounded sample of the same BED
    e2e = None
    if args.no_extras:
        if rank == 0:
            emit(json.dumps({'profile_run': True, 'kernel_ms': kern_ms, 'ms_per_step': total_ms / args.steps}))
        if world > 1:
            dist.destroy_process_group()
        return
    sample = min(bed.numel(), args.e2e_bytes)
    cut = int((bed[:sample] == 10).nonzero()[-1].item()) + 1
    hbuf = torch.empty(cut, dtype=torch.uint8).pin_memory()
    hbuf.copy_(bed[:cut])
    sample_lines = int((bed[:cut] == 10).sum().item())
    stage = torch.empty(2 * (64 << 20), dtype=torch.uint8, device=dev)
    hres = torch.empty(4, dtype=torch.int64).pin_memory()

    hlinks = {}

    def e2e_step():
        eng.table.zero_()
        eng.ingest_host(hbuf, stage=stage)
        ok, oc, m = eng._compact()            # (key, count) list of the non-empty cells, key-sorted
        if 'k' not in hlinks or hlinks['k'].numel() < m:
            hlinks['k'] = torch.empty(m, dtype=torch.int64).pin_memory()
            hlinks['c'] = torch.empty(m, dtype=torch.int64).pin_memory()
        hlinks['k'][:m].copy_(ok[:m], non_blocking=True)      # D2H of the step's result
        hlinks['c'][:m].copy_(oc[:m], non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return m * 16 + 8

    for _ in range(2):
        d2h = e2e_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    w0 = time.perf_counter()
    reps = max(2, min(args.steps, 5))
    for _ in range(reps):
        d2h = e2e_step()
    torch.cuda.synchronize()
    e2e_s = (time.perf_counter() - w0) / reps
    if world > 1:
        t = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t[0])
    del hres
    # the ceiling of that leg: the same pinned bytes through plain cudaMemcpyAsync in the same 64 MiB pieces
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    half = stage.numel() // 2
    ev0.record()
    for o in range(0, cut, half):
        k = min(half, cut - o)
        stage[:k].copy_(hbuf[o:o + k], non_blocking=True)
    ev1.record()
    torch.cuda.synchronize()
    h2d_gbs = cut / (ev0.elapsed_time(ev1) * 1e-3) / 1e9
    e2e_pairs = sample_lines / (nlines / pairs)
    e2e = {'value': e2e_pairs * world / e2e_s, 'unit': UNIT, 'h2d_bytes_per_step': int(cut), 'd2h_bytes_per_step': int(d2h),
           'gbs': cut / e2e_s / 1e9, 'h2d_copy_only_gbs': h2d_gbs,
           'sample': 'first %d bytes of the same BED from pinned host memory via salsa_bed_ingest_dense_host '
                     '(64 MiB double-buffered chunks) + D2H of the compacted link list' % cut}

    # the north-star multi-GPU split on BASELINE configs[4]'s shape, and a small oracle-checked run of the same route
    nbytes_main, table_cells = bed.numel(), n * n * 4
    sparse = None
    merge_ok = None
    if not args.no_sparse:
        bed_prefix = bed[:min(bed.numel(), args.cpu_sample_bytes)].clone() if (world == 1) else None
        del bed, expected
        eng.table = None
        torch.cuda.empty_cache()
        merge_ok = small_merge_check(eng, dev, rank, world)
        sparse = sparse_leg(args, eng, dev, rank, world)
        torch.cuda.empty_cache()
    else:
        bed_prefix = bed
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    peak, which = peaks()
    bed_bytes = nbytes_main
    achieved = bed_bytes / (kern_ms * 1e-3) / 1e9
    out = {
        'metric': METRIC, 'value': pairs * world / (total_ms / args.steps * 1e-3), 'unit': UNIT, 'n_gpus': world,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': total_ms / args.steps, 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'u8', 'data': 'synthetic',
        'config': {'workload': 'BASELINE configs[1]: 3.1 Gb assembly, 3,000 contigs, %d Hi-C pairs per GPU (%d BED lines, %d bytes of '
                               '6-column BED text resident in HBM); dense link table %d cells; -e GATC,GANTC only affects scoring'
                               % (pairs, nlines, bed_bytes, table_cells),
                   'l2_policy': 'input (%.1f GB) larger than L2' % (bed_bytes / 1e9),
                   'parallelism': 'read pairs sharded over %d GPU(s); NCCL all-reduce of the dense table' % world,
                   'parity_full_size': 'table == generator expected table: %s; counted pairs %d' % (parity, stt.pairs)},
        'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
                     'traffic': (ncu_traffic_ratio() * bed_bytes) if ncu_traffic_ratio() else None,
                     'traffic_source': 'ncu --set full dram read+write of a 4.70 GB launch (profiles/r02_v8_ingest_metrics.txt), '
                                       'scaled to this launch by algorithmic bytes',
                     'peak_source': which + ' (MEASURED_PEAKS.json hbm_gbs, burst copy)',
                     'kernel': 'bedstream::bed_stream_kernel<LinkEmit<dense>>', 'algorithmic_bytes_per_launch': bed_bytes,
                     'kernel_ms': kern_ms,
                     'atomics_per_s': stt.pairs / (args.steps + args.warmup) / (kern_ms * 1e-3),
                     'atomics_note': 'one 64-bit RED into the dense table per counted pair (L2); ~1 % of the ~2e11/s that '
                                     'spread-address REDG sustains chip-wide (B300_MICROARCH.md: 1.29 cycles/lane/SM), so the '
                                     'histogram is nowhere near an atomic limit and is not privatised in shared memory'},
        'e2e': e2e,
        'gpu_launches': args.steps * 2,      # per step: the sample kernel that picks the special-byte test + the scan kernel
        'clocks': clocks,
    }
    if sparse is not None:
        out['sparse'] = sparse
        out['sparse_merge_vs_oracle_small_case'] = bool(merge_ok)
    if world == 1:
        out['cpu_baseline'] = cpu_baseline(bed_prefix, asm, nlines / pairs, args.cpu_sample_bytes, threads=1, eng=eng)
        if not args.no_iteration:
            del bed_prefix
            torch.cuda.empty_cache()
            out['iteration'] = iteration_leg(args)
    emit(json.dumps(out))


def iteration_leg(args):
    """The second half of BASELINE.json's metric: seconds per pipeline iteration at configs[1]'s assembly (3.1 Gb, 3,000
    contigs) on a `--iteration-pairs` BED, through the drop-in run_pipeline.py (tools/bench_pipeline.py: files on disk in,
    files out, the step programs in one process with the BED resident in HBM), plus the CPU path's seconds for the hot
    steps on a stated sample of the same BED."""
    cmd = [sys.executable, os.path.join(ROOT, 'tools', 'bench_pipeline.py'), str(args.iteration_pairs), '--cfg2', '--gpu-only',
           '--cpu-sample', str(args.iteration_cpu_lines)]
    t0 = time.time()
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        return {'error': r.stderr[-1500:]}
    rec = json.loads(r.stdout.strip().splitlines()[-1])
    rec['leg_wall_s'] = round(time.time() - t0, 1)
    return rec


def cpu_baseline(bed, asm, lines_per_pair, sample_bytes, threads, eng=None):
    """oracle/ C restatement of make_links.py's loop, timed on this box's host cores
    on a bounded prefix of the same BED (the one place bench.py executes oracle/).
    The same prefix also goes through the CUDA path, and the two link tables must be
    equal: every bench run is a parity run against the oracle as well."""
    import oracle
    oracle.build()
    sample = min(bed.numel(), sample_bytes)
    cut = int((bed[:sample] == 10).nonzero()[-1].item()) + 1
    hb = bytes(bed[:cut].cpu().numpy())
    n = asm.ncontig
    one = np.ones(n, np.int64)
    t0 = time.perf_counter()
    res = oracle.make_links(hb, asm.names, asm.lengths.astype(float), np.ones(n, np.uint8), one, one, nthreads=threads)
    dt = time.perf_counter() - t0
    lines = hb.count(b'\n')
    out = {'value': lines / lines_per_pair / dt, 'unit': UNIT, 'cores': threads, 'kind': 'port',
           'sample': 'first %d bytes (%d lines) of the same BED, oracle/salsa_oracle.c oracle_make_links, %d thread(s), %.2f s'
                     % (cut, lines, threads, dt), 'links': int(res['count'].sum())}
    if eng is not None:
        eng.set_contigs(asm.names, asm.lengths, one, one)
        eng.begin_links('dense')
        eng.ingest_device(bed[:cut])
        stt = eng.stats_read()
        f, s2, e, c = eng.links()
        same = (len(c) == len(res['count']) and (f == res['first']).all() and (s2 == res['second']).all()
                and (e == res['ends']).all() and (c == res['count']).all() and stt.pairs == res['npairs_counted'])
        if not same:
            raise SystemExit('PARITY FAILURE: CUDA link table of the CPU-baseline prefix differs from the oracle')
        out['gpu_parity_on_sample'] = 'CUDA link list of the same %d bytes == oracle link list (%d links, %d counted pairs): True' % (
            cut, len(c), stt.pairs)
    return out


def run_reference(args):
    """The reference's own CPU implementation of the path.  make_links.py is
    Python-2 only and cannot run in this image, so this arm times the oracle's C
    restatement of its loop on all host threads (kind 'port')."""
    rank = int(os.environ.get('RANK', 0))
    if rank != 0:
        return
    import oracle
    from salsa_b200 import synth
    oracle.build()
    cfg = dict(synth.CONFIGS[args.config])
    asm = synth.make_assembly(seed=20261106 + int(args.config[-1]), **cfg)
    n = asm.ncontig
    sample_pairs = args.ref_pairs
    rec = synth.gen_records(asm, sample_pairs, seed=1000, device='cpu')
    bed = bytes(synth.format_bed(rec).numpy())
    nlines = rec['ctg'].numel()
    threads = os.cpu_count() or 1
    one = np.ones(n, np.int64)

    def step():
        return oracle.make_links(bed, asm.names, asm.lengths.astype(float), np.ones(n, np.uint8), one, one, nthreads=threads)

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        res = step()
    dt = (time.perf_counter() - t0) / args.steps
    val = sample_pairs / dt
    out = {'impl': 'reference', 'metric': METRIC, 'value': val, 'unit': UNIT, 'n_gpus': int(os.environ.get('WORLD_SIZE', 1)),
           'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': dt * 1e3, 'higher_is_better': True, 'scaling': 'weak',
           'vs_baseline': None, 'dtype': 'u8', 'data': 'synthetic',
           'config': {'workload': 'BASELINE configs[1] shape (3.1 Gb assembly, 3,000 contigs); each step = a bounded sample of '
                                  '%d pairs (%d BED lines, %d bytes) of that workload' % (sample_pairs, nlines, len(bed))},
           'cpu_baseline': {'value': val, 'unit': UNIT, 'cores': threads, 'kind': 'port',
                            'sample': '%d pairs per step; oracle/salsa_oracle.c (C restatement of make_links.py:60-146), '
                                      'byte-range sharded over %d threads' % (sample_pairs, threads)},
           'e2e': {'value': val, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
           'gpu_launches': 0, 'links': int(res['count'].sum())}
    emit(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--config', default='cfg2')
    ap.add_argument('--pairs', type=int, default=None, help='pairs per GPU (default: the config\'s)')
    ap.add_argument('--e2e-bytes', type=int, default=4 << 30)
    ap.add_argument('--cpu-sample-bytes', type=int, default=1 << 30)
    ap.add_argument('--no-extras', action='store_true', help='timed region only (for ncu runs): no e2e / cpu legs')
    ap.add_argument('--ref-pairs', type=int, default=10_000_000)
    ap.add_argument('--no-sparse', action='store_true', help='skip the configs[4]-shape sparse leg')
    ap.add_argument('--sparse-pairs', type=int, default=500_000_000, help='pairs per GPU of the sparse leg')
    ap.add_argument('--no-iteration', action='store_true', help='skip the seconds-per-pipeline-iteration leg (N = 1 only)')
    ap.add_argument('--iteration-pairs', type=int, default=100_000_000)
    ap.add_argument('--iteration-cpu-lines', type=int, default=4_000_000)
    args = ap.parse_args()
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    args.warmup = max(args.warmup, 3) if args.impl == 'ours' else args.warmup
    if args.pairs is None:
        from salsa_b200 import synth
        args.pairs = synth.CONFIGS[args.config]['pairs']
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
