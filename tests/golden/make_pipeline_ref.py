"""Regenerates tests/golden/pipeline_ref/: a two-iteration run of the reference's OWN programs on a
small synthetic data set, stage by stage, so that every drop-in step can be fed the reference's
inputs and compared with the reference's outputs.

Run in the build container:   python tests/golden/make_pipeline_ref.py

Python steps: /root/reference/*.py executed by ref_py2emu.py (Python-2 semantics restored, logic =
the reference's source text).  break_contigs: oracle/_ref (the reference's C++ compiled unmodified).
`sort -k 5 -gr` under LC_ALL=C as run_pipeline.py does.  The misjoin report of stage B is written by
hand (the real one is empty: a 0.5 Mb toy genome has no 4 Mb windows) so that the scaffold splitting
and the avoid list are exercised.

Files (gzip, mtime 0): the inputs, and per stage the files the reference wrote.  Pickles are stored
as JSON lists of [name, path] in dict order.
"""
import gzip
import json
import os
import pickle
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
sys.path.insert(0, HERE)

import oracle  # noqa: E402
import pipeline_util as pu  # noqa: E402
import ref_py2emu as R  # noqa: E402

OUT = os.path.join(HERE, 'pipeline_ref')
PARAMS = dict(ncontig=36, genome=500_000, len_lo=2_000, len_hi=30_000, npairs=14_000, seed=4242, nchrom=3)


def put(name, blob):
    with open(os.path.join(OUT, name + '.gz'), 'wb') as raw:
        with gzip.GzipFile(fileobj=raw, mode='wb', mtime=0) as g:
            g.write(blob)


def put_file(name, path):
    put(name, open(path, 'rb').read())


def put_pickle(name, path):
    with open(path, 'rb') as f:
        d = pickle.load(f)
    put(name, json.dumps([[k, v] for k, v in d.items()]).encode())


def ref(script, *argv):
    out, rc = R.run(script, list(argv))
    assert rc == 0, (script, argv, rc)
    return out


def main():
    oracle.build()
    shutil.rmtree(OUT, ignore_errors=True)
    os.makedirs(OUT)
    w = tempfile.mkdtemp()
    data, d = w + '/data', w + '/out'
    pu.write_dataset(data, **PARAMS)
    os.makedirs(d)
    for fn in ('assembly.fa', 'contigs.len', 'reads.bed'):
        put_file('in.' + fn, data + '/' + fn)
    shutil.copy(data + '/contigs.len', d + '/scaffold_length_iteration_1')
    shutil.copy(data + '/reads.bed', d + '/alignment_iteration_1.bed')
    shutil.copy(data + '/assembly.fa', d + '/assembly.cleaned.fasta')
    env = dict(os.environ, LC_ALL='C')

    def links(it):
        ref('make_links.py', '-b', '%s/alignment_iteration_%d.bed' % (d, it), '-d', d, '-i', it)
        ref('fast_scaled_scores.py', '-d', d, '-i', it)
        with open('%s/contig_links_scaled_sorted_iteration_%d' % (d, it), 'wb') as f:
            subprocess.run(['sort', '-k', '5', '-gr', '%s/contig_links_scaled_iteration_%d' % (d, it)], stdout=f, env=env, check=True)
        for fn in ('contig_links_iteration_%d', 'contig_links_scaled_iteration_%d', 'contig_links_scaled_sorted_iteration_%d'):
            put_file('s%d.' % it + fn % it, d + '/' + fn % it)

    def layout(it):
        ref('layout_unitigs.py', '-x', 'abc', '-l', '%s/contig_links_scaled_sorted_iteration_%d' % (d, it), '-c', '1000', '-i', it, '-d', d)
        put_pickle('s%d.layout.scaffolds_iteration_%d.json' % (it, it), '%s/scaffolds_iteration_%d.p' % (d, it))
        for fn in ('breakpoints_iteration_%d.txt', 'alignment_iteration_%d.bed', 'scaffold_length_iteration_%d', 're_counts_iteration_%d'):
            put_file('s%d.layout.' % it + fn % (it + 1), d + '/' + fn % (it + 1))

    def misjoin(it, report):
        r = subprocess.run([oracle.REF_DIR + '/break_contigs', '-a', '%s/alignment_iteration_%d.bed' % (d, it + 1), '-b',
                            '%s/breakpoints_iteration_%d.txt' % (d, it + 1), '-l', '%s/scaffold_length_iteration_%d' % (d, it + 1),
                            '-i', str(it + 1), '-s', '100'], capture_output=True, check=True)
        put('s%d.break_contigs.stdout' % it, r.stdout)
        open('%s/misasm_iteration_%d.report' % (d, it + 1), 'w').write(report)
        put('s%d.misasm_iteration_%d.report' % (it, it + 1), report.encode())
        ref('refactor_breaks.py', '-d', d, '-i', it + 1)
        put_pickle('s%d.refactor.scaffolds_iteration_%d.json' % (it, it), '%s/scaffolds_iteration_%d.p' % (d, it))
        for fn in ('links_avoid_iteration_%d', 'alignment_iteration_%d.bed', 'scaffold_length_iteration_%d', 're_counts_iteration_%d'):
            put_file('s%d.refactor.' % it + fn % (it + 1), d + '/' + fn % (it + 1))

    # stage 0 (-m yes): correct.py with a hand-written input_breaks (three contigs cut; side outputs only,
    # the stages below run on the uncut assembly)
    c = w + '/clean'
    os.makedirs(c)
    lens = [ln.split() for ln in open(data + '/contigs.len')]
    cuts = ''.join('%s\t%d\n' % (a[0], int(a[1]) * frac // 100) for a, frac in zip([lens[1], lens[7], lens[20]], (50, 25, 80)))
    open(c + '/input_breaks', 'w').write(cuts)
    ref('correct.py', data + '/assembly.fa', c + '/input_breaks', data + '/reads.bed', c)
    put('s0.input_breaks', cuts.encode())
    put_file('s0.alignment_iteration_1.tmp.bed', c + '/alignment_iteration_1.tmp.bed')
    put_file('s0.asm.cleaned.fasta', c + '/asm.cleaned.fasta')
    # correct.py only writes its BED buffer every 1,000,000 lines and never flushes the rest (correct.py:86-91):
    # the BED above comes out empty.  Same step on the BED repeated 40 times (> 1 M lines); only the digest is kept.
    import hashlib
    open(c + '/big.bed', 'wb').write(open(data + '/reads.bed', 'rb').read() * 40)
    assert R.run('correct.py', [data + '/assembly.fa', c + '/input_breaks', c + '/big.bed', c], as_function=True)[1] == 0
    big = open(c + '/alignment_iteration_1.tmp.bed', 'rb').read()
    put('s0.big40.digest', json.dumps(dict(sha256=hashlib.sha256(big).hexdigest(), bytes=len(big), lines=big.count(b'\n'))).encode())
    put('s1.re_counts_iteration_1', ref('RE_sites.py', '-a', d + '/assembly.cleaned.fasta', '-e', 'GATC').encode())
    open(d + '/re_counts_iteration_1', 'w').write(gzip.open(OUT + '/s1.re_counts_iteration_1.gz', 'rt').read())
    links(1)
    layout(1)
    # flag the 2nd and 4th join of the first long scaffold and the 1st join of the second one
    bp = [ln.split() for ln in open(d + '/breakpoints_iteration_2.txt')]
    longs = [a for a in bp if len(a) >= 6]
    report = '%s\t%s\t%s\n' % (longs[0][0], longs[0][2], longs[0][4])
    if len(longs) > 1:
        report += '%s\t%s\n' % (longs[1][0], longs[1][1])
    report += 'Total Joins = 9\nSuspicious Joins = 3\nPercent Suspicious Joins = 33.3333\n'
    misjoin(1, report)
    links(2)
    layout(2)
    misjoin(2, open(d + '/misasm_iteration_3.report', 'w') and 'Total Joins = 0\nSuspicious Joins = 0\nPercent Suspicious Joins = -nan\n')
    ref('get_seq.py', '-a', d + '/assembly.cleaned.fasta', '-f', d + '/scaffolds_FINAL.fasta', '-g', d + '/scaffolds_FINAL.agp', '-p',
        d + '/scaffolds_iteration_2.p')
    put_file('final.scaffolds_FINAL.agp', d + '/scaffolds_FINAL.agp')
    put_file('final.scaffolds_FINAL.fasta', d + '/scaffolds_FINAL.fasta')
    json.dump(dict(params=PARAMS, how='tests/golden/make_pipeline_ref.py'), open(OUT + '/README.json', 'w'), indent=1, sort_keys=True)
    print('pipeline_ref written:', sum(os.path.getsize(os.path.join(OUT, f)) for f in os.listdir(OUT)), 'bytes')


if __name__ == '__main__':
    main()
