#!/usr/bin/env python
"""Wall seconds per pipeline iteration (the second half of BASELINE.json's metric) on
BASELINE configs[0]: synthetic 2 Mb genome in 200 contigs, 1 M Hi-C pairs, -e GATC -i 3.

Runs the drop-in run_pipeline.py twice over the same inputs: with the CUDA step programs
(salsa_b200/bin) and with the CPU steps (oracle C port + compiled reference coverage tools; the
reference's own Python-2 steps cannot run in this image; the BED re-projection is the oracle's
Python twin).  The greedy layout, misjoin splitting and AGP writer are the same code in both runs.
Prints one JSON line.
"""
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))

import pipeline_util as pu  # noqa: E402


def run(bin_dir, data, out, extra):
    cmd = [sys.executable, os.path.join(pu.BIN, 'run_pipeline.py'), '--bin', bin_dir, '-a', data + '/assembly.fa', '-l',
           data + '/contigs.len', '-b', data + '/reads.bed', '-o', out] + extra
    t0 = time.time()
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    wall = time.time() - t0
    steps, per_iter = {}, []
    for ln in open(out + '/timings.tsv'):
        k, v = ln.split('\t')
        if '--verbose' in sys.argv:
            sys.stderr.write('  %-24s %8.3f s\n' % (k, float(v)))
        steps[k] = steps.get(k, 0.0) + float(v)
        if k == 'make_links.py':                      # every iteration opens with it (run_pipeline.py:271-385)
            per_iter.append({})
        if per_iter and k in ITER_STEPS:
            per_iter[-1][k] = per_iter[-1].get(k, 0.0) + float(v)
    iters = len([f for f in os.listdir(out) if f.startswith('contig_links_iteration_')])
    run.per_iteration = per_iter
    return wall, steps, iters


ITER_STEPS = ('make_links.py', 'fast_scaled_scores.py', 'layout_unitigs.py', 'break_contigs', 'refactor_breaks.py')


def cpu_sample(data, gpu_out, work, nlines):
    """The CPU path's seconds for the hot steps of ONE iteration on the first `nlines` lines of the same BED: the
    oracle's C port of make_links + fast_scaled_scores (1 thread) and the compiled reference break_contigs_start."""
    import oracle
    oracle.build()
    obin = pu.make_oracle_bin(work + '/obin')
    os.makedirs(work + '/cpu', exist_ok=True)
    sample = work + '/sample.bed'
    with open(data + '/reads.bed', 'rb') as f, open(sample, 'wb') as g:
        for i, ln in enumerate(f):
            if i >= nlines:
                break
            g.write(ln)
    for nm in ('scaffold_length_iteration_1', 're_counts_iteration_1'):
        os.symlink(os.path.join(gpu_out, nm), os.path.join(work, 'cpu', nm))
    # break_contigs_start's cost is the genome (difference arrays + 11 Kadane passes), not the BED: a tenth of the contigs
    with open(data + '/contigs.len', 'rb') as f:
        lens = f.read().splitlines(True)
    nlen = max(1, len(lens) // 10)
    open(work + '/sample.len', 'wb').write(b''.join(lens[:nlen]))
    out = {'sample': 'first %d lines of the same BED; oracle C port of make_links / fast_scaled_scores (1 thread); compiled reference '
                     'break_contigs_start on the first %d of %d contigs (its cost is the genome, not the BED)' % (nlines, nlen, len(lens)),
           'lines': nlines, 'contigs_for_break_contigs_start': nlen}
    for key, cmd in (('make_links.py', [sys.executable, obin + '/make_links.py', '-b', sample, '-d', work + '/cpu', '-i', '1']),
                     ('fast_scaled_scores.py', [sys.executable, obin + '/fast_scaled_scores.py', '-d', work + '/cpu', '-i', '1']),
                     ('break_contigs_start', [obin + '/break_contigs_start', '-a', sample, '-l', work + '/sample.len', '-s', '100'])):
        t0 = time.time()
        r = subprocess.run(cmd, capture_output=True)
        assert r.returncode == 0, r.stderr[-2000:]
        out[key] = round(time.time() - t0, 3)
    return out


def main():
    pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    d = tempfile.mkdtemp(dir=os.environ.get('SALSA_BENCH_TMP'))
    data = d + '/data'
    if '--cfg2' in sys.argv:
        # BASELINE configs[1]'s assembly: 3.1 Gb in 3,000 contigs, -e GATC,GANTC (the BED holds `pairs` pairs)
        t0 = time.time()
        pu.write_dataset(data, ncontig=3000, genome=3_100_000_000, len_lo=50_000, len_hi=20_000_000, npairs=pairs, seed=20261108, nchrom=23)
        sys.stderr.write('data set written in %.1f s\n' % (time.time() - t0))
        args = ['-e', 'GATC,GANTC', '-i', '3', '-m', 'yes' if '--misasm' in sys.argv else 'no']
        shape = 'BASELINE configs[1] assembly: 3.1 Gb, 3,000 contigs, %d pairs, %s' % (pairs, ' '.join(args))
    else:
        pu.write_dataset(data, ncontig=200, genome=2_000_000, len_lo=2_000, len_hi=40_000, npairs=pairs, seed=20261107, nchrom=4)
        args = ['-e', 'GATC', '-i', '3', '-m', 'no']
        shape = 'BASELINE configs[0]: 2 Mb genome, 200 contigs, %d pairs, %s' % (pairs, ' '.join(args))
    wg, sg, ig = run(pu.BIN, data, d + '/gpu', args)
    if '--gpu-only' in sys.argv:
        per = run.per_iteration
        rec = {'metric': 'seconds_per_pipeline_iteration', 'config': shape, 'pairs': pairs, 'iterations_run': ig,
               'first_iteration_s': round(sum(per[0].values()), 3) if per else None,
               'mean_iteration_s': round(sum(sum(p.values()) for p in per) / max(1, len(per)), 3),
               'per_iteration_steps_s': [{k: round(v, 3) for k, v in p.items()} for p in per],
               'gpu_path': {'wall_s': wg, 'startup_s': sg.get('startup', 0.0), 'steps_s': sg},
               'note': 'one iteration = make_links -> fast_scaled_scores (+ sort) -> layout_unitigs -> break_contigs -> '
                       'refactor_breaks (run_pipeline.py:271-385); the first one reads the whole input BED, later ones the '
                       're-projected BED, which shrinks as contigs join'}
        if '--cpu-sample' in sys.argv:
            rec['cpu_path'] = cpu_sample(data, d + '/gpu', d, int(sys.argv[sys.argv.index('--cpu-sample') + 1]))
        print(json.dumps(rec))
        if '--keep' not in sys.argv:
            import shutil
            shutil.rmtree(d, ignore_errors=True)
        return
    import oracle
    oracle.build()
    obin = pu.make_oracle_bin(d + '/obin')
    wo, so, io = run(obin, data, d + '/cpu', args)
    same = open(d + '/gpu/scaffolds_FINAL.agp').read() == open(d + '/cpu/scaffolds_FINAL.agp').read()
    hot = ('make_links.py', 'fast_scaled_scores.py', 'RE_sites.py', 'break_contigs', 'break_contigs_start')
    print(json.dumps({
        'metric': 'seconds_per_pipeline_iteration', 'config': shape,
        'iterations_run': ig, 'identical_final_agp': same,
        'gpu_path': {'wall_s': wg, 's_per_iteration': wg / ig, 'startup_s': sg.get('startup', 0.0),
                     's_per_iteration_after_startup': (sum(sg.values()) - sg.get('startup', 0.0)) / ig,
                     'hot_steps_s': sum(sg.get(k, 0) for k in hot), 'steps_s': sg},
        'cpu_path': {'wall_s': wo, 's_per_iteration': wo / io, 'hot_steps_s': sum(so.get(k, 0) for k in hot), 'steps_s': so,
                     'kind': 'oracle C port + compiled reference break_contigs*, 1 core'},
        'note': 'GPU path: the Python steps run inside run_pipeline.py (one CUDA context; "startup" = torch import + context '
                'creation + library load, once per run), break_contigs* are calls into the library.  CPU path: '
                'every step is a process, the BED re-projection inside layout_unitigs.py / refactor_breaks.py is the oracle twin '
                '(pure Python).  The greedy layout, misjoin splitting and AGP writer are the same host code in both.'}))


if __name__ == '__main__':
    main()
