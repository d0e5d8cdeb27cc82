#!/usr/bin/env python3
"""Drop-in for the reference's run_pipeline.py: same flags, same files in the -o directory, same
resume-by-file rules, same iteration loop (including its off-by-one: `-i K` with K >= 2 computes
K + 1 iterations and emits scaffolds_iteration_K.p as FINAL, run_pipeline.py:256-269,377-385; an
unchanged NG50 stops the loop and emits the previous iteration, :361-374).

Steps are separate programs found next to this file (or in --bin), exactly as in the reference:
the GPU-backed make_links.py / fast_scaled_scores.py / RE_sites.py / break_contigs_start /
break_contigs, the host-side correct.py / correct_links / layout_unitigs.py / refactor_breaks.py /
get_seq.py, and coreutils `sort -k 5 -gr` (LC_ALL=C is forced so the tie order is defined).
"""
import argparse
import os
import stat
import subprocess
import sys


def first_ratio_below_one(path):
    """check() of run_pipeline.py:9-16: looks at the first sorted link only."""
    with open(path, 'r') as f:
        for line in f:
            return float(line.split()[4]) < 1
    return None


def nonempty(path):
    return os.path.isfile(path) and os.path.getsize(path) > 0


def ng50_of(lengths, genome_size=0):
    total = genome_size or sum(lengths.values())
    acc = 0
    for ln in sorted(lengths.values(), reverse=True):
        acc += ln
        if acc >= total // 2:
            return ln
    return None


def _flush_written_behind():
    """BED files a step left to a background writer must be complete before anyone outside this process reads them."""
    mod = sys.modules.get('salsa_b200.reproject')
    if mod is not None:
        mod.flush()


def main(argv=None):
    try:
        return _main(argv)
    finally:
        _flush_written_behind()


def _main(argv=None):
    here = os.path.dirname(os.path.abspath(__file__))
    ap = argparse.ArgumentParser(description="SALSA Iterative Pipeline")
    ap.add_argument('-a', '--assembly', required=True, help='Path to initial assembly')
    ap.add_argument('-l', '--length', required=True, help='Length of contigs at start')
    ap.add_argument('-b', '--bed', required=True, help='Bed file of alignments sorted by read names')
    ap.add_argument('-o', '--output', default='SALSA_output', help='Output directory to put results')
    ap.add_argument('-O', '--output-original-coords', action='store_true', dest='orig_coords',
                    help='Also write AGP files in the coordinates of the input assembly (requires -m yes)')
    ap.add_argument('-c', '--cutoff', default=1000, help='Minimum contig length to scaffold, default=1000')
    ap.add_argument('-g', '--gfa', default='abc', help='GFA file for assembly')
    ap.add_argument('-e', '--enzyme', default='AAGCTT', help='Restriction Enzyme used for experiment')
    ap.add_argument('-i', '--iter', default=3, help='Number of iterations to run, default = 3')
    ap.add_argument('-x', '--dup', default='abc', help='File containing duplicated contig information')
    ap.add_argument('-s', '--exp', default=0, help='Expected Genome size of the assembled genome')
    ap.add_argument('-m', '--clean', default='yes', help='"yes" to find misassemblies in the input assembly')
    ap.add_argument('-f', '--filter', default='no', help='Filter bed file for contigs present in the assembly')
    ap.add_argument('-p', '--prnt', default='no', help='"yes" to output scaffolds and agp for each iteration')
    ap.add_argument('--bin', default=here, help=argparse.SUPPRESS)      # where the step programs live
    ap.add_argument('--subprocess', action='store_true', help=argparse.SUPPRESS)   # spawn every step like the reference
    args = ap.parse_args(argv)
    if args.orig_coords and args.clean != 'yes':
        print("ERROR: You must use -m,--clean 'yes' with -O,--output-original-coords.", file=sys.stderr)
        return 1
    B, out = args.bin, args.output
    py = sys.executable
    env = dict(os.environ, LC_ALL='C')
    genome_size = int(args.exp)
    os.makedirs(out, exist_ok=True)
    log = open(out + '/commands.log', 'w', 1)

    import time
    timings = open(out + '/timings.tsv', 'a', 1)      # extra to the reference: wall seconds per step

    # The Python steps of THIS package run inside this process by default: one interpreter, one CUDA
    # context and one loaded library for the whole run instead of ~4 s of start-up per step.  The
    # command lines are still logged as the reference would run them; --subprocess restores spawning.
    inproc = (os.path.abspath(B) == here) and not args.subprocess
    modules = {}

    def run_inprocess(words, redirect):
        import contextlib
        import importlib.util
        import io
        path = words[1]
        name = os.path.basename(path)
        if name not in modules:
            if here not in sys.path:
                sys.path.insert(0, here)
            spec = importlib.util.spec_from_file_location('salsa_step_' + name[:-3], path)
            modules[name] = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(modules[name])
        buf = io.StringIO()
        try:
            with contextlib.redirect_stdout(buf):
                rc = modules[name].main(words[2:]) or 0
        except SystemExit as e:
            rc = e.code if isinstance(e.code, int) else (0 if e.code is None else 1)
        except Exception as e:                       # what would be a non-zero exit of the step program
            print('%s: %s: %s' % (name, type(e).__name__, e), file=sys.stderr)
            rc = 1
        if redirect and rc == 0:
            with open(redirect, 'w') as f:
                f.write(buf.getvalue())
        elif redirect:
            open(redirect, 'w').close()
        return rc

    def run_coverage_inprocess(step, words, redirect):
        """break_contigs_start / break_contigs through the library call their main() wraps: same CUDA
        context as the other steps, and the BED straight from HBM when a step left it there."""
        import ctypes
        if here not in sys.path:
            sys.path.insert(0, here)
        import _common
        from salsa_b200 import _lib, reproject
        eng = _common.engine()
        argv = words[1:]
        bed, ptr, nbytes = None, None, 0
        if '-a' in argv:
            try:
                bed = reproject.resident_bed(argv[argv.index('-a') + 1], eng.device)
            except (MemoryError, OSError):
                bed = None                                   # too large to keep resident / missing: the call reads the file
        if bed is not None and bed.numel():
            ptr, nbytes = bed.data_ptr(), bed.numel()
        else:
            reproject.flush()                                # the program is going to read the file itself
        arr = (ctypes.c_char_p * len(argv))(*[a.encode() for a in argv])
        sys.stdout.flush()
        return _lib.lib().salsa_break_contigs_run(int(step == 'break_contigs'), len(argv), arr, ptr, nbytes,
                                                  redirect.encode() if redirect else None)

    def sh(cmd, fatal=True, cleanup=None):
        print(cmd)
        log.write(cmd + '\n')
        t0 = time.time()
        words = cmd.split()
        step = os.path.basename(words[1] if words[0] == py else words[0])
        redirect = None
        if inproc and not modules.get('started') and (step.endswith('.py') or step.startswith('break_contigs')):
            # one-off: torch import, CUDA context, library load -- timed on its own line, not inside the first step
            if here not in sys.path:
                sys.path.insert(0, here)
            import _common
            _common.engine()
            _common.KEEP_RESIDENT = True      # BEDs stay in HBM between steps; their files are written behind (flushed at exit)
            modules['started'] = True
            timings.write('startup\t%.3f\n' % (time.time() - t0))
            t0 = time.time()
        if inproc and '>' in words and (words[0] == py or step in ('break_contigs', 'break_contigs_start')):
            redirect = words[words.index('>') + 1]
            words = words[:words.index('>')]
        if inproc and words[0] == py and step.endswith('.py'):
            prof_dir = os.environ.get('SALSA_PROFILE_STEPS')       # development aid: cProfile of every in-process step
            if prof_dir:
                import cProfile
                import io
                import pstats
                pr = cProfile.Profile()
                pr.enable()
            rc, errtext = run_inprocess(words, redirect), ''
            if prof_dir:
                pr.disable()
                buf = io.StringIO()
                pstats.Stats(pr, stream=buf).sort_stats('cumtime').print_stats(28)
                modules['nprof'] = modules.get('nprof', 0) + 1
                open(os.path.join(prof_dir, '%02d_%s.txt' % (modules['nprof'], step)), 'w').write(buf.getvalue())
        elif inproc and step in ('break_contigs', 'break_contigs_start'):
            rc, errtext = run_coverage_inprocess(step, words, redirect), ''
        else:
            if '.bed' in cmd:
                _flush_written_behind()
            r = subprocess.run(cmd, shell=True, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
            rc, errtext = r.returncode, r.stdout + r.stderr
        timings.write('%s\t%.3f\n' % (step, time.time() - t0))
        if rc != 0:
            print(errtext, file=sys.stderr)
            if cleanup and os.path.isfile(cleanup):
                os.remove(cleanup)
            if fatal:
                sys.exit(1)
        return rc == 0

    def get_seq(pickle_iter, tag):
        cmd = '%s %s/get_seq.py -a %s/assembly.cleaned.fasta -f %s/scaffolds_%s.fasta -g %s/scaffolds_%s.agp -p %s/scaffolds_iteration_%s.p' % (
            py, B, out, out, tag, out, tag, out, pickle_iter)
        if args.orig_coords:
            cmd += ' -b %s/input_breaks -G %s/scaffolds_%s.original-coordinates.agp' % (out, out, tag)
        return cmd

    if not os.path.exists(out + '/scaffold_length_iteration_1'):
        sh('cp %s %s/scaffold_length_iteration_1' % (args.length, out))
        os.chmod(out + '/scaffold_length_iteration_1', stat.S_IRUSR | stat.S_IWUSR | stat.S_IRGRP | stat.S_IWGRP | stat.S_IROTH)
    bed1 = out + '/alignment_iteration_1.bed'
    if not nonempty(bed1):
        if os.path.lexists(bed1):
            os.remove(bed1)
        if args.filter == 'yes':
            os.system("cut -f 1 %s | grep -v '>'  > %s/contig_names.txt" % (args.length, out))
            os.system("grep -f %s/contig_names.txt -w  %s > %s" % (out, args.bed, bed1))
        else:
            os.symlink(os.path.abspath(args.bed), bed1)
    if not os.path.lexists(out + '/assembly.cleaned.fasta'):
        os.symlink(os.path.abspath(args.assembly), out + '/assembly.cleaned.fasta')

    if args.clean == 'yes':                                           # run_pipeline.py:120-136 (failures are not fatal there)
        sh('%s/break_contigs_start -a %s -l %s/scaffold_length_iteration_1 -s 100 > %s/input_breaks' % (B, bed1, out, out), fatal=False)
        sh('%s %s/correct.py  %s %s/input_breaks %s %s' % (py, B, args.assembly, out, bed1, out), fatal=False)
        os.system('mv %s/alignment_iteration_1.tmp.bed %s' % (out, bed1))
        os.system('mv %s/asm.cleaned.fasta %s/assembly.cleaned.fasta' % (out, out))

    it = 1
    if not nonempty('%s/re_counts_iteration_%d' % (out, it)):
        sh('%s %s/RE_sites.py -a %s/assembly.cleaned.fasta -e %s > %s/re_counts_iteration_%d' % (py, B, out, args.enzyme, out, it))

    ng50 = []

    def link_steps(it, first):
        print('Starting Iteration %d' % it, file=sys.stderr)
        links = '%s/contig_links_iteration_%d' % (out, it)
        if not nonempty(links):
            extra = ' -x ' + args.dup if first else ''
            sh('%s %s/make_links.py -b %s/alignment_iteration_%d.bed -d %s -i %d%s' % (py, B, out, it, out, it, extra), cleanup=links)
        scaled = '%s/contig_links_scaled_iteration_%d' % (out, it)
        if not (nonempty(scaled) if first else os.path.isfile(scaled)):
            sh('%s %s/fast_scaled_scores.py -d %s -i %d' % (py, B, out, it), cleanup=scaled)
        srt = '%s/contig_links_scaled_sorted_iteration_%d' % (out, it)
        if not nonempty(srt):
            sh('sort -k 5 -gr %s > %s' % (scaled, srt), cleanup=srt)
        return srt

    def misjoin_steps(it, first):
        rep = '%s/misasm_iteration_%d.report' % (out, it + 1)
        if not (nonempty(rep) if first else os.path.isfile(rep)):
            sh('%s/break_contigs -a %s/alignment_iteration_%d.bed -b %s/breakpoints_iteration_%d.txt -l %s/scaffold_length_iteration_%d '
               '-i %d -s 100 > %s' % (B, out, it + 1, out, it + 1, out, it + 1, it + 1, rep))
        done = '%s/misasm_%d.DONE' % (out, it + 1)
        if not (nonempty(done) if first else os.path.isfile(done)):
            tail = '' if first else ' > %s/misasm_%d.log' % (out, it + 1)
            sh('%s %s/refactor_breaks.py -d %s -i %d%s' % (py, B, out, it + 1, tail))
        if args.prnt == 'yes':
            sh(get_seq(it, 'ITERATION_%d' % it), fatal=False)

    def lengths_of(it):
        d = {}
        with open('%s/scaffold_length_iteration_%d' % (out, it)) as f:
            for line in f:
                a = line.split()
                d[a[0]] = int(a[1])
        return d

    # ---- iteration 1 (run_pipeline.py:143-269)
    srt = link_steps(1, True)
    if args.gfa != 'abc' and not os.path.isfile(out + '/tmp.links'):
        sh('%s/correct_links -g %s -l %s > %s/tmp.links' % (B, args.gfa, srt, out))
        os.system('mv %s/tmp.links %s' % (out, srt))
    if not os.path.isfile(out + '/scaffolds_iteration_1.p'):
        sh('%s %s/layout_unitigs.py -x %s -l %s -c %s -i 1 -d %s' % (py, B, args.gfa, srt, args.cutoff, out))
    misjoin_steps(1, True)
    it = 2
    ng50.append(ng50_of(lengths_of(it), genome_size))
    if it - 1 == int(args.iter):
        sh(get_seq(args.iter, 'FINAL'), fatal=False)
        return 0
    # ---- iterations >= 2 (run_pipeline.py:271-385)
    while True:
        srt = link_steps(it, False)
        if first_ratio_below_one(srt):
            print('WARNING: Not enough Hi-C reads for scaffolding', file=sys.stderr)
        if not nonempty('%s/scaffolds_iteration_%d.p' % (out, it)):
            sh('%s %s/layout_unitigs.py -x abc -l %s -c %s -i %d -d %s' % (py, B, srt, args.cutoff, it, out))
        misjoin_steps(it, False)
        ng50.append(ng50_of(lengths_of(it + 1), genome_size))
        if ng50[-1] == ng50[-2]:
            sh(get_seq(it - 1, 'FINAL'), fatal=False)
            return 0
        if it - 1 == int(args.iter):
            sh(get_seq(args.iter, 'FINAL'), fatal=False)
            return 0
        it += 1


if __name__ == '__main__':
    sys.exit(main())
