#!/usr/bin/env python
"""Per-kernel measurements at BASELINE configs[1]/[2] sizes for the rows of SURVEY section 8 that
bench.py (the headline ingest metric) does not cover: RE-site counting, coverage, scoring.

    python tools/bench_kernels.py [--quick]  > profiles/rNN_kernels.jsonl

One JSON line per kernel group: CUDA-event time (best of `reps` after a warm-up), algorithmic
bytes, achieved GB/s and the fraction of the measured HBM peak, plus a size-independent parity
property checked at full size.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from salsa_b200 import coverage, synth  # noqa: E402
from salsa_b200.engine import Engine  # noqa: E402
from salsa_b200.re_sites import count_sites  # noqa: E402


def peak():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    return json.load(open(p))['hbm_gbs'] if os.path.exists(p) else 6650.0


def emit(name, ms, alg_bytes, **kw):
    gbs = alg_bytes / (ms * 1e-3) / 1e9
    print(json.dumps(dict(kernel=name, ms=ms, algorithmic_bytes=int(alg_bytes), achieved_gbs=gbs, hbm_peak_gbs=peak(),
                          frac=gbs / peak(), **kw)), flush=True)


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


def bench_re(eng, asm, dev):
    """3.1 Gb FASTA text built in HBM (60-column lines), -e GATC,GANTC (Arima)."""
    lens = (asm.lengths // 60) * 60
    G = int(lens.sum())
    g = torch.Generator(device=dev)
    g.manual_seed(5)
    codes = torch.randint(0, 4, (G,), generator=g, device=dev, dtype=torch.uint8)
    letters = 65 + 2 * (codes == 1).to(torch.uint8) + 6 * (codes == 2).to(torch.uint8) + 19 * (codes == 3).to(torch.uint8)  # A C G T
    rows = letters.view(-1, 60)
    rows = torch.cat([rows, torch.full((rows.shape[0], 1), 10, dtype=torch.uint8, device=dev)], 1).view(-1)
    pieces, pos = [], 0
    for i, ln in enumerate(lens):
        nb = int(ln) // 60 * 61
        pieces.append(torch.tensor(list(b'>' + asm.names[i] + b' len=%d\n' % ln), dtype=torch.uint8, device=dev))
        pieces.append(rows[pos:pos + nb])
        pos += nb
    fasta = torch.cat(pieces)
    del rows, pieces
    for enz in ('GATC,GANTC',):
        best = {}
        for rep in range(3):
            tm = {}
            names, left, right, seqlen = count_sites(eng, fasta, enz, timers=tm)
            torch.cuda.synchronize()
            for k, ev in tm.items():
                if k != '_sizes':
                    best[k] = min(best.get(k, 1e30), ev[0].elapsed_time(ev[1]))
            sizes = tm['_sizes']
        # parity property at full size: GATC + GA[ACGT]TC occurrences counted straight from the base codes
        G_, A_, T_, C_ = 2, 0, 3, 1
        c = codes
        m4 = (c[:-3] == G_) & (c[1:-2] == A_) & (c[2:-1] == T_) & (c[3:] == C_)
        m5 = (c[:-4] == G_) & (c[1:-3] == A_) & (c[3:-1] == T_) & (c[4:] == C_)
        ok = True
        off = np.concatenate([[0], np.cumsum(lens)])
        for i in list(range(0, len(lens), 97)):
            a, b = int(off[i]), int(off[i + 1])
            half = (b - a) // 2
            l4 = int(m4[a:a + half].sum()) if half <= b - a - 3 else 0
            t4 = int(m4[a:b - 3].sum())
            l5 = int(m5[a:a + half].sum())
            t5 = int(m5[a:b - 4].sum())
            ok &= (int(left[i]) == l4 + l5) and (int(right[i]) == (t4 - l4) + (t5 - l5))
        assert ok, 'RE parity failure'
        assert (seqlen == lens).all()
        nf = sizes['fasta_bytes']
        emit('fasta_headers', best['headers'], nf, bases=sizes['bases'], note='FASTA text read once')
        emit('fasta_count+scan+span_lengths', best['count'], nf, note='FASTA text read once (kept-byte counts per 4 KiB tile)')
        emit('fasta_pack', best['pack'], nf + sizes['plane_bytes'], note='FASTA text read once + 3 bit planes written (3G/8 bytes)')
        emit('re_count[%s]' % enz, best['re_count'], sizes['plane_bytes'], patterns=sizes['npatterns'],
             note='bit planes read once for all %d expanded patterns' % sizes['npatterns'], parity='sampled contigs match direct base-code count')
    del fasta, codes


def bench_cov(eng, asm, dev, pairs):
    """break_contigs_start at cfg3 shape: BED scan -> diff arrays for >= 1 Mb contigs, prefix + 11 Kadane windows."""
    n = asm.ncontig
    lens = asm.lengths
    covoff, ndiff = coverage.layout(lens)
    eng.set_contigs(asm.names, lens)
    buf = torch.empty(int(pairs * 2.02 * 47) + (1 << 20), dtype=torch.uint8, device=dev)
    pos, done = 0, 0
    while done < pairs:
        m = min(20_000_000, pairs - done)
        rec = synth.gen_records(asm, m, seed=900 + done, device=dev, read_id_base=done)
        out = synth.format_bed(rec, out=buf[pos:])
        pos += out.numel()
        done += m
        del rec, out
    bed = buf[:pos]
    diff = torch.zeros(ndiff + 1, dtype=torch.int32, device=dev)
    touched = torch.zeros(n + 1, dtype=torch.uint8, device=dev)
    import ctypes as C
    from salsa_b200._lib import check
    check(eng.L.salsa_ingest_stats_reset(eng.stats.data_ptr(), None))

    def run_scan():
        diff.zero_()
        coverage.scan(eng, bed, lens, covoff, ndiff, diff=diff, touched=touched)
    ms = timed(run_scan)
    st = eng.stats_read()
    emit('cov_scan_start (BED -> diff arrays)', ms, bed.numel(), pairs=pairs, counted_pairs_per_pass=int(st.pairs // 4),
         note='BED text read once; +2 int32 atomics per accepted same-contig pair; includes zeroing %d MB of diff arrays' % (ndiff * 4 >> 20))
    # ---- BED re-projection (update_bed pass) on the same BED: 3,000 contigs placed on 24 scaffolds, 10 % unplaced
    from salsa_b200 import reproject
    rng = np.random.default_rng(5)
    nsc = 24
    sc_of = rng.integers(0, nsc, n).astype(np.int32)
    sc_of[rng.random(n) < 0.1] = -1
    offs = np.zeros(n, dtype=np.int64)
    acc = np.zeros(nsc, dtype=np.int64)
    for i in range(n):
        if sc_of[i] >= 0:
            offs[i] = acc[sc_of[i]]
            acc[sc_of[i]] += int(lens[i])
    rev = (rng.random(n) < 0.5).astype(np.uint8)
    snames = [b'scaffold_%d' % (k + 1) for k in range(nsc)]
    for refactor in (False, True):
        res = {}

        def run_rp():
            res['out'] = reproject.reproject(eng, bed, refactor, sc_of, offs, np.asarray(lens, np.int64), rev, snames)
        ms = timed(run_rp)
        out = res['out']
        nl = int((out == 10).sum())
        first = out.clone()
        run_rp()
        assert nl % 2 == 0 and torch.equal(first, res['out'])
        emit('bed_reproject (%s variant: scan + sort + emit x2)' % ('refactor' if refactor else 'layout'), ms,
             bed.numel() + out.numel(), pairs=pairs, kept_pairs=nl // 2, out_bytes=int(out.numel()),
             note='BED read once by the candidate scan, candidate lines re-read by the two emit passes, output written once; '
                  'includes two host round trips (candidate count, output size)', parity='deterministic; 2 lines per kept pair')
        del out, first, res
    # parity property: every contig's differences sum to zero, coverage never negative
    ids = [i for i in range(n) if covoff[i] != coverage.NOSEED and touched[i]]
    segs = []
    for i in ids:
        sz = int(lens[i]) + 10
        segs.append(coverage.Seg(int(covoff[i]), int(covoff[i]), coverage.NOSEED, sz, sz // 100, 9 * sz // 10))
    total = sum(s.len for s in segs)
    keep = diff.clone()

    def run_an():
        diff.copy_(keep)
        return coverage.analyse(eng, diff, diff, segs)
    ms_all = timed(run_an)
    ms_copy = timed(lambda: diff.copy_(keep))
    hs, ks, ke = run_an()
    assert int(diff.min()) >= 0
    ends = torch.tensor([int(covoff[i]) + int(lens[i]) + 9 for i in ids], device=dev)
    assert int(diff[ends].abs().sum()) == 0            # coverage returns to zero at the end of every contig
    emit('cov_analyse (prefix + 11-threshold Kadane)', ms_all - ms_copy, total * 4 * 4, segments=len(segs), elements=total,
         note='4 B/element: diff read (sums), diff read + coverage written (scan), coverage read once for all 11 thresholds')
    del buf, diff


def bench_avoid(eng, asm, dev, pairs):
    """make_links.py of an iteration >= 2: the same scan with links_avoid_iteration_N loaded (the tile scanner with the
    walk-back / replay of bed_scan.cuh).  Three avoid lists: a handful of scaffold pairs (what refactor_breaks.py writes),
    10 % of the contigs involved, every contig involved (the worst case: no line is certainly kept)."""
    n = asm.ncontig
    one = np.ones(n, np.int64)
    buf = torch.empty(int(pairs * 2.02 * 47) + (1 << 20), dtype=torch.uint8, device=dev)
    pos, done = 0, 0
    while done < pairs:
        m = min(20_000_000, pairs - done)
        rec = synth.gen_records(asm, m, seed=700 + done, device=dev, read_id_base=done)
        out = synth.format_bed(rec, out=buf[pos:])
        pos += out.numel()
        done += m
        del rec, out
    bed = buf[:pos]
    rng = np.random.default_rng(11)
    counted = {}
    for label, k in (('no avoid list (streaming scanner)', 0), ('30 avoid pairs', 30), ('avoid pairs over 10 % of the contigs', n // 20),
                     ('every contig in an avoid pair', n)):
        if k == 0:
            avoid = []
        elif k == n:
            perm = rng.permutation(n)
            avoid = [(asm.names[perm[i]], asm.names[perm[(i + 1) % n]]) for i in range(n)]
        else:
            idx = rng.choice(n, size=2 * k, replace=False)
            avoid = [(asm.names[idx[2 * i]], asm.names[idx[2 * i + 1]]) for i in range(k)]
        eng.set_contigs(asm.names, asm.lengths, re_left=one, re_right=one, avoid=avoid)
        eng.begin_links('dense')

        def run():
            eng.ingest_device(bed)
        ms = timed(run)
        st = eng.stats_read()
        counted[label] = int(st.pairs)
        emit('bed_ingest, %s' % label, ms, bed.numel(), pairs=pairs, avoid_pairs=len(avoid),
             counted_pairs_total=int(st.pairs), note='dense table; BED text read once')
    eng.set_contigs(asm.names, asm.lengths)
    del buf, bed


def bench_score(eng, dev, n=3000, E=18_000_000):
    rng = np.random.default_rng(1)
    a = rng.integers(0, n, E * 2)
    b = rng.integers(0, n, E * 2)
    m = a < b
    keys = np.unique(((a[m] * n + b[m]) << 2) | rng.integers(0, 4, int(m.sum())))[:E]
    E = len(keys)
    counts = rng.integers(1, 5000, E)
    dk = torch.from_numpy(keys.astype(np.int64)).to(dev)
    dc = torch.from_numpy(counts.astype(np.int64)).to(dev)
    rl = torch.from_numpy(rng.integers(1, 4000, n)).to(dev)
    rr = torch.from_numpy(rng.integers(1, 4000, n)).to(dev)
    f64 = lambda: torch.empty(E, dtype=torch.float64, device=dev)      # noqa: E731
    i64 = lambda: torch.empty(E, dtype=torch.int64, device=dev)        # noqa: E731
    i32 = lambda: torch.empty(E, dtype=torch.int32, device=dev)        # noqa: E731
    u8 = lambda: torch.empty(E, dtype=torch.uint8, device=dev)         # noqa: E731
    score, w12, dig, ex, z0, u, v = f64(), f64(), i64(), i32(), u8(), i32(), i32()
    from salsa_b200._lib import check
    L = eng.L
    ms = timed(lambda: check(L.salsa_links_score(dk.data_ptr(), dc.data_ptr(), E, n, rl.data_ptr(), rr.data_ptr(), score.data_ptr(),
                                                 z0.data_ptr(), dig.data_ptr(), ex.data_ptr(), w12.data_ptr(), u.data_ptr(),
                                                 v.data_ptr(), None)))
    emit('links_score (+ exact 12-digit round trip)', ms, E * (16 + 37), edges=E,
         note='16 B/edge read, 37 B/edge written; exact %.12g digits from 128-bit integers (1e-8 <= x < 1e12), 320-bit route beyond')
    txt = ['%.12g' % x for x in score[:200000].cpu().numpy()]
    assert (w12[:200000].cpu().numpy() == np.array([float(t) for t in txt])).all()
    scratch = torch.empty(L.salsa_scaled_scratch_bytes(2 * n) // 8 + 1, dtype=torch.int64, device=dev)
    best, ratio, rd, rk, bd, bk, i1 = f64(), f64(), i64(), i32(), i64(), i32(), u8()
    ms = timed(lambda: check(L.salsa_scaled_scores(u.data_ptr(), v.data_ptr(), w12.data_ptr(), E, 2 * n, scratch.data_ptr(),
                                                   best.data_ptr(), i1.data_ptr(), ratio.data_ptr(), rd.data_ptr(), rk.data_ptr(),
                                                   bd.data_ptr(), bk.data_ptr(), None)))
    emit('scaled_scores (node top-2 + ratio + 2 round trips)', ms, E * (16 * 3 + 41), edges=E,
         note='16 B/edge read three times, 41 B/edge written; u side reduced per warp before the atomics')
    nm = torch.zeros(2 * n, dtype=torch.float64, device=dev)
    nm.scatter_reduce_(0, u.long(), w12, 'amax')
    nm.scatter_reduce_(0, v.long(), w12, 'amax')
    assert bool((best <= torch.maximum(nm[u.long()], nm[v.long()])).all())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--quick', action='store_true')
    ap.add_argument('--only', default='')
    args = ap.parse_args()
    dev = torch.device('cuda', 0)
    eng = Engine(0)
    cfg = dict(synth.CONFIGS['cfg2'])
    if args.quick:
        cfg.update(ncontig=300, genome=310_000_000)
    asm = synth.make_assembly(seed=20261108, **cfg)
    if args.only in ('', 're'):
        bench_re(eng, asm, dev)
        torch.cuda.empty_cache()
    if args.only in ('', 'cov'):
        bench_cov(eng, asm, dev, 10_000_000 if args.quick else 100_000_000)
        torch.cuda.empty_cache()
    if args.only in ('', 'avoid'):
        bench_avoid(eng, asm, dev, 10_000_000 if args.quick else 100_000_000)
        torch.cuda.empty_cache()
    if args.only in ('', 'score'):
        bench_score(eng, dev, E=2_000_000 if args.quick else 18_000_000)


if __name__ == '__main__':
    main()
