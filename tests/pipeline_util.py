"""Helpers for the end-to-end pipeline tests: a synthetic data set on disk, and an
"oracle bin" directory whose accelerated steps (make_links.py, fast_scaled_scores.py,
RE_sites.py, break_contigs, break_contigs_start, the BED pass inside layout_unitigs.py and
refactor_breaks.py) are the CPU oracle / the compiled reference, while the host-side steps are
the product's own -- so a run with it and a run with salsa_b200/bin differ ONLY in the hot-path
implementation."""
import os
import stat
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, 'salsa_b200', 'bin')

ORACLE_STEP = '''#!%(py)s
import sys
sys.path.insert(0, %(root)r)
import argparse
import oracle
%(body)s
'''

BODIES = {
    'make_links.py': '''
p = argparse.ArgumentParser()
p.add_argument('-b'); p.add_argument('-i'); p.add_argument('-d'); p.add_argument('-x')
a = p.parse_args()
lines = oracle.make_links_files(a.b, a.d, a.i)
print(len(lines))
''',
    'fast_scaled_scores.py': '''
p = argparse.ArgumentParser()
p.add_argument('-i'); p.add_argument('-d')
a = p.parse_args()
oracle.fast_scaled_scores_files(a.d, a.i)
''',
    'RE_sites.py': '''
p = argparse.ArgumentParser()
p.add_argument('-a'); p.add_argument('-e')
a = p.parse_args()
try:
    names, left, right, seqlen = oracle.re_sites(open(a.a, 'rb').read(), a.e)
except oracle.OracleError as e:
    sys.stderr.write(str(e) + '\\n')
    sys.exit(1)
sys.stdout.write(''.join('%s %d %d\\n' % (n.decode('latin-1'), l, r) for n, l, r in zip(names, left, right)))
''',
}

# the product's host-side step with the oracle's BED pass injected in place of the CUDA one
HOST_STEP = '''#!%(py)s
import sys
sys.path.insert(0, %(root)r)
sys.path.insert(0, %(bin)r)
import importlib.util
from oracle import py_twin
spec = importlib.util.spec_from_file_location('product_step', %(bin)r + '/%(name)s')
mod = importlib.util.module_from_spec(spec)
spec.loader.exec_module(mod)


def main(argv=None):
    return mod.main(argv, bed_pass=py_twin.reproject_bed)


if __name__ == '__main__':
    sys.exit(main())
'''


def make_oracle_bin(dst):
    os.makedirs(dst, exist_ok=True)
    for name, body in BODIES.items():
        p = os.path.join(dst, name)
        open(p, 'w').write(ORACLE_STEP % dict(py=sys.executable, root=ROOT, body=body))
    ref = os.path.join(ROOT, 'oracle', '_ref')
    for prog in ('break_contigs', 'break_contigs_start'):
        p = os.path.join(dst, prog)
        open(p, 'w').write('#!/bin/sh\nexec %s/%s "$@"\n' % (ref, prog))
        os.chmod(p, os.stat(p).st_mode | stat.S_IXUSR | stat.S_IXGRP | stat.S_IXOTH)
    for name in ('layout_unitigs.py', 'refactor_breaks.py'):
        p = os.path.join(dst, name)
        if os.path.lexists(p):
            os.remove(p)
        open(p, 'w').write(HOST_STEP % dict(py=sys.executable, root=ROOT, bin=BIN, name=name))
    for name in ('correct.py', 'get_seq.py', '_common.py', 'correct_links'):
        d = os.path.join(dst, name)
        if os.path.lexists(d):
            os.remove(d)
        os.symlink(os.path.join(BIN, name), d)
    return dst


def write_dataset(d, ncontig, genome, len_lo, len_hi, npairs, seed, nchrom=3, with_gfa=False):
    """assembly.fa, contigs.len (fai-like), reads.bed (+ assembly.gfa) under d."""
    from salsa_b200 import synth
    asm = synth.make_assembly(ncontig, genome, len_lo, len_hi, nchrom=nchrom, seed=seed)
    os.makedirs(d, exist_ok=True)
    open(d + '/assembly.fa', 'wb').write(synth.gen_fasta(asm))
    open(d + '/contigs.len', 'wb').write(asm.lengths_text())
    with open(d + '/reads.bed', 'wb') as f:
        if npairs <= 20_000_000:
            rec = synth.gen_records(asm, npairs, seed + 1, noise=0.03, triple=0.002)
            f.write(bytes(synth.format_bed(rec).numpy()))
        else:                                # large benchmark inputs: generate on the GPU, 20 M pairs at a time
            for done in range(0, npairs, 20_000_000):
                m = min(20_000_000, npairs - done)
                rec = synth.gen_records(asm, m, seed + 1 + done, device='cuda', read_id_base=done, noise=0.03, triple=0.002)
                synth.format_bed(rec).cpu().numpy().tofile(f)
                del rec
    if with_gfa:
        rng = np.random.default_rng(seed)
        lines = ['S\t%s\t*\tLN:i:%d' % (n.decode(), l) for n, l in zip(asm.names, asm.lengths)]
        for k in range(asm.ncontig - 1):
            if asm.chrom_of_ordered[k] == asm.chrom_of_ordered[k + 1] and rng.random() < 0.3:
                a, b = asm.order[k], asm.order[k + 1]
                lines.append('L\t%s\t%s\t%s\t%s\t0M' % (asm.names[a].decode(), '-' if asm.orient[a] else '+',
                                                        asm.names[b].decode(), '-' if asm.orient[b] else '+'))
        open(d + '/assembly.gfa', 'w').write('\n'.join(lines) + '\n')
    return asm


def comparable(outdir):
    """file name -> content made order-insensitive where the reference's own order is hash order."""
    res = {}
    for fn in sorted(os.listdir(outdir)):
        p = os.path.join(outdir, fn)
        if os.path.islink(p) or not os.path.isfile(p) or fn.endswith('.p') or fn in ('commands.log', 'timings.tsv') or fn.endswith('.log'):
            continue
        data = open(p, 'rb').read()
        if fn.startswith(('misasm_iteration_', 'input_breaks', 're_counts_iteration_1')):
            data = b'\n'.join(sorted(data.split(b'\n')))
        res[fn] = data
    return res
