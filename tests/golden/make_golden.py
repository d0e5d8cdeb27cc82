"""Regenerates tests/golden/*.json.

Run in the build container (where /root/reference exists and oracle/_ref has
been compiled by oracle/Makefile):   python tests/golden/make_golden.py

coverage_ref.json : stdout of the UNMODIFIED reference programs
                    (oracle/_ref/break_contigs_start, break_contigs) on the
                    deterministic cases in tests/cases.py.
links_twin.json   : the reference's Python steps cannot run here (Python 2
                    only), so these vectors come from the line-by-line py3
                    twin (oracle/py_twin.py) -- they pin the C oracle and the
                    CUDA path against a second, independent restatement.
"""
import json
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))

import cases  # noqa: E402
import oracle  # noqa: E402
from oracle import py_twin  # noqa: E402
from salsa_b200 import synth  # noqa: E402


def main():
    oracle.build()
    d = tempfile.mkdtemp()
    cov = {'start': {}, 'joins': {}}
    for nm, (seed, lens, holes, npairs) in cases.COV_START_CASES.items():
        names, L, bed = cases.coverage_case(seed, lens, holes, npairs)
        open(d + '/b.bed', 'wb').write(bed)
        open(d + '/l.txt', 'wb').write(cases.lengths_text(names, L))
        r = subprocess.run([oracle.REF_DIR + '/break_contigs_start', '-a', d + '/b.bed', '-l', d + '/l.txt', '-s', '100'],
                           capture_output=True, check=True)
        cov['start'][nm] = sorted(r.stdout.decode().splitlines())
    for nm, (seed, lens, holes, npairs, bp) in cases.COV_JOIN_CASES.items():
        names, L, bed = cases.coverage_case(seed, lens, holes, npairs, scaffold_style=True)
        bpn = {names[k]: v for k, v in bp.items()}
        open(d + '/b.bed', 'wb').write(bed)
        open(d + '/l.txt', 'wb').write(cases.lengths_text(names, L))
        open(d + '/bp.txt', 'wb').write(cases.breakpoints_text(bpn))
        r = subprocess.run([oracle.REF_DIR + '/break_contigs', '-a', d + '/b.bed', '-l', d + '/l.txt', '-b', d + '/bp.txt',
                            '-i', '2', '-s', '100'], capture_output=True, check=True)
        cov['joins'][nm] = sorted(r.stdout.decode().splitlines())
    json.dump(cov, open(os.path.join(HERE, 'coverage_ref.json'), 'w'), indent=1, sort_keys=True)

    # links / RE / scaled vectors from the py3 twin on a small synthetic case
    asm, rec, bed = synth.small_case(ncontig=12, genome=200000, npairs=3000, seed=1)
    fa = synth.gen_fasta(asm)
    open(d + '/a.fa', 'wb').write(fa)
    # the inputs themselves are committed (gzip, mtime 0) so the vectors do not
    # depend on torch's CPU RNG staying stable
    import gzip
    for fn, blob in (('links_case.bed.gz', bed), ('links_case.fa.gz', fa), ('links_case.len.gz', asm.lengths_text())):
        with open(os.path.join(HERE, fn), 'wb') as raw:
            with gzip.GzipFile(fileobj=raw, mode='wb', mtime=0) as g:
                g.write(blob)
    out = {'params': dict(ncontig=12, genome=200000, npairs=3000, seed=1)}
    for enz in ('GATC', 'GATC,GANTC', 'DNASE', 'AAGCTT'):
        out['re_' + enz] = py_twin.re_sites(d + '/a.fa', enz)
    open(d + '/scaffold_length_iteration_1', 'wb').write(asm.lengths_text())
    open(d + '/re_counts_iteration_1', 'w').write('\n'.join(out['re_GATC,GANTC']) + '\n')
    open(d + '/b.bed', 'wb').write(bed)
    out['links'] = sorted(py_twin.make_links(d + '/b.bed', d, 1))
    out['scaled'] = sorted(py_twin.fast_scaled_scores(d, 1))
    # iteration 2 flavour: an avoid list (stale prev_line rule)
    for f in ('scaffold_length', 're_counts'):
        os.link(d + '/%s_iteration_1' % f, d + '/%s_iteration_2' % f)
    open(d + '/links_avoid_iteration_2', 'w').write('ctg000001\tctg000002\nctg000003\tctg000007\nctg000000\tctg000005\n')
    out['links_avoid'] = sorted(py_twin.make_links(d + '/b.bed', d, 2))
    out['avoid'] = [['ctg000001', 'ctg000002'], ['ctg000003', 'ctg000007'], ['ctg000000', 'ctg000005']]
    json.dump(out, open(os.path.join(HERE, 'links_twin.json'), 'w'), indent=1, sort_keys=True)
    print('golden written')


if __name__ == '__main__':
    main()
