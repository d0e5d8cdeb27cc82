"""Regenerates tests/golden/*.json.

Run in the build container (where /root/reference exists and oracle/_ref has
been compiled by oracle/Makefile):   python tests/golden/make_golden.py

coverage_ref.json : stdout of the UNMODIFIED reference programs
                    (oracle/_ref/break_contigs_start, break_contigs) on the
                    deterministic cases in tests/cases.py.
correct_links_ref.json : stdout of the UNMODIFIED reference correct_links.cpp,
                    compiled against the test-only boost stand-in
                    oracle/_shim (boost itself is not in the image), on
                    cases.CORRECT_LINKS_CASES.
links_twin.json   : vectors from the line-by-line py3 twin (oracle/py_twin.py),
                    a second, independent restatement.
links_ref.json    : the same vectors plus a messy-text case, produced by the
                    reference's OWN Python-2 scripts executed from
                    /root/reference by ref_py2emu.py (Python-2 semantics
                    restored, program logic = the reference's source text).
                    (+ messy_case.*.gz, its inputs.)
make_pipeline_ref.py does the same for a whole two-iteration run.
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

    # correct_links: stdout of the UNMODIFIED reference source compiled against the boost stand-in of oracle/_shim
    cl = {}
    for nm, mk in cases.CORRECT_LINKS_CASES.items():
        gfa, links = mk()
        open(d + '/g.gfa', 'w').write(gfa)
        open(d + '/hl.txt', 'w').write(links)
        r = subprocess.run([oracle.REF_DIR + '/correct_links', '-g', d + '/g.gfa', '-l', d + '/hl.txt'], capture_output=True, check=True)
        cl[nm] = r.stdout.decode().split('\n')[:-1]
    json.dump(cl, open(os.path.join(HERE, 'correct_links_ref.json'), 'w'), indent=1, sort_keys=True)

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
    json.dump(reference_vectors(d, out, bed, fa, asm.lengths_text()), open(os.path.join(HERE, 'links_ref.json'), 'w'),
              indent=1, sort_keys=True)
    print('golden written')


def reference_vectors(d, twin, bed, fa, lens_text):
    """links_ref.json: the same vectors as links_twin.json, plus a messy-text case, produced by the
    reference's OWN make_links.py / fast_scaled_scores.py / RE_sites.py executed from /root/reference
    under the Python-2 emulation of ref_py2emu.py (read its header for what is emulated)."""
    import gzip
    import numpy as np
    import ref_py2emu as R

    def lines(path):
        return sorted(open(path).read().splitlines(keepends=True))

    ref = {'params': twin['params'], 'avoid': twin['avoid'], 'how': 'tests/golden/ref_py2emu.py over /root/reference/*.py'}
    for enz in ('GATC', 'GATC,GANTC', 'DNASE', 'AAGCTT'):
        text, rc = R.run('RE_sites.py', ['-a', d + '/a.fa', '-e', enz])
        assert rc == 0
        ref['re_' + enz] = text.splitlines()
    r = tempfile.mkdtemp()
    open(r + '/scaffold_length_iteration_1', 'wb').write(lens_text)
    open(r + '/re_counts_iteration_1', 'w').write('\n'.join(ref['re_GATC,GANTC']) + '\n')
    open(r + '/b.bed', 'wb').write(bed)
    text, rc = R.run('make_links.py', ['-b', r + '/b.bed', '-d', r, '-i', '1'])
    assert rc == 0
    ref['links'], ref['links_stdout'] = lines(r + '/contig_links_iteration_1'), text
    assert R.run('fast_scaled_scores.py', ['-d', r, '-i', '1'])[1] == 0
    ref['scaled'] = lines(r + '/contig_links_scaled_iteration_1')
    for f in ('scaffold_length', 're_counts'):
        os.link(r + '/%s_iteration_1' % f, r + '/%s_iteration_2' % f)
    open(r + '/links_avoid_iteration_2', 'w').write(''.join('%s\t%s\n' % tuple(p) for p in twin['avoid']))
    assert R.run('make_links.py', ['-b', r + '/b.bed', '-d', r, '-i', '2'])[1] == 0
    ref['links_avoid'] = lines(r + '/contig_links_iteration_2')
    assert R.run('fast_scaled_scores.py', ['-d', r, '-i', '2'])[1] == 0
    ref['scaled_avoid'] = lines(r + '/contig_links_scaled_iteration_2')
    # messy text: whitespace variety, long names, signs, '\r', several '/', no final newline
    rng = np.random.default_rng(77)
    names = cases.MESSY_NAMES
    L = [int(rng.integers(2000, 6000)) for _ in names]
    mbed = cases.messy_bed(rng, names, 4000, L)
    m = tempfile.mkdtemp()
    mlen = b''.join(b'%s\t%d\n' % (nm, ln) for nm, ln in zip(names, L))
    mre = b''.join(b'%s\t%d\t%d\n' % (nm, 3 + 2 * i, 40 - 3 * i) for i, nm in enumerate(names))
    open(m + '/scaffold_length_iteration_1', 'wb').write(mlen)
    open(m + '/re_counts_iteration_1', 'wb').write(mre)
    open(m + '/b.bed', 'wb').write(mbed)
    assert R.run('make_links.py', ['-b', m + '/b.bed', '-d', m, '-i', '1'])[1] == 0
    assert R.run('fast_scaled_scores.py', ['-d', m, '-i', '1'])[1] == 0
    ref['messy_links'] = lines(m + '/contig_links_iteration_1')
    ref['messy_scaled'] = lines(m + '/contig_links_scaled_iteration_1')
    for fn, blob in (('messy_case.bed.gz', mbed), ('messy_case.len.gz', mlen), ('messy_case.re.gz', mre)):
        with open(os.path.join(HERE, fn), 'wb') as raw:
            with gzip.GzipFile(fileobj=raw, mode='wb', mtime=0) as g:
                g.write(blob)
    return ref


if __name__ == '__main__':
    main()
