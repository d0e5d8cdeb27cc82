"""Every step fed the REFERENCE's inputs must write the REFERENCE's outputs.

tests/golden/pipeline_ref/ holds a two-iteration run of the reference's own programs (its Python
scripts executed by tests/golden/ref_py2emu.py, its C++ break_contigs compiled unmodified), stage
by stage; tests/golden/make_pipeline_ref.py generated it.  Here the CPU side is checked: the oracle
(C port and py3 twin) for the hot steps, and the product's host code (layout, re-projection tables,
misjoin splitting, AGP writer) with the oracle's BED pass injected.  tests/test_gpu_reproject.py and
tests/test_gpu_pipeline.py repeat the hot steps on the GPU against the same files.
"""
import gzip
import json
import os
import pickle
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, 'tests', 'golden', 'pipeline_ref')
sys.path.insert(0, os.path.join(ROOT, 'salsa_b200', 'bin'))


def get(name):
    return gzip.open(os.path.join(REF, name + '.gz'), 'rb').read()


def scaffolds(name):
    return {k: v for k, v in json.loads(get(name))}


def stage(tmp_path, files):
    d = str(tmp_path)
    for dst, src in files.items():
        if src.endswith('.json'):
            with open(os.path.join(d, dst), 'wb') as f:
                pickle.dump(scaffolds(src), f)
        else:
            open(os.path.join(d, dst), 'wb').write(get(src))
    return d


def canonical(scaf):
    return sorted(min(tuple(p), tuple(reversed(p))) for p in scaf.values())


def lengths():
    return {a.split()[0].decode(): int(a.split()[1]) for a in get('in.contigs.len').splitlines()}


BASE = {'scaffold_length_iteration_1': 'in.contigs.len', 're_counts_iteration_1': 's1.re_counts_iteration_1',
        'alignment_iteration_1.bed': 'in.reads.bed'}


@pytest.mark.parametrize('impl', ['c', 'twin'])
def test_hot_steps(oracle_lib, tmp_path, impl):
    from oracle import py_twin
    o = oracle_lib
    names, left, right, _ = o.re_sites(get('in.assembly.fa'), 'GATC')
    assert ''.join('%s %d %d\n' % (n.decode(), a, b) for n, a, b in zip(names, left, right)).encode() == get('s1.re_counts_iteration_1')
    d = stage(tmp_path, dict(BASE, **{'scaffold_length_iteration_2': 's1.refactor.scaffold_length_iteration_2',
                                      're_counts_iteration_2': 's1.refactor.re_counts_iteration_2',
                                      'links_avoid_iteration_2': 's1.refactor.links_avoid_iteration_2',
                                      'alignment_iteration_2.bed': 's1.refactor.alignment_iteration_2.bed'}))
    for it in (1, 2):
        bed = '%s/alignment_iteration_%d.bed' % (d, it)
        if impl == 'c':
            got = o.make_links_files(bed, d, it)
        else:
            got = py_twin.make_links(bed, d, it)
            open('%s/contig_links_iteration_%d' % (d, it), 'w').write(''.join(got))
        want = get('s%d.contig_links_iteration_%d' % (it, it)).decode().splitlines(keepends=True)
        assert sorted(got) == sorted(want) and len(want) > 20
        # the next program reads the reference's file (its line order decides nothing, but keep it exact)
        open('%s/contig_links_iteration_%d' % (d, it), 'wb').write(get('s%d.contig_links_iteration_%d' % (it, it)))
        got = o.fast_scaled_scores_files(d, it) if impl == 'c' else py_twin.fast_scaled_scores(d, it)
        assert ''.join(got).encode() == get('s%d.contig_links_scaled_iteration_%d' % (it, it))
        open(d + '/scaled', 'w').write(''.join(got))
        r = subprocess.run(['sort', '-k', '5', '-gr', d + '/scaled'], capture_output=True, env=dict(os.environ, LC_ALL='C'))
        assert r.stdout == get('s%d.contig_links_scaled_sorted_iteration_%d' % (it, it))


@pytest.mark.parametrize('it', [1, 2])
def test_update_bed_after_layout(tmp_path, it):
    from oracle import py_twin
    from salsa_b200.host import reproject
    d = stage(tmp_path, BASE)
    reproject.update_bed(d, scaffolds('s%d.layout.scaffolds_iteration_%d.json' % (it, it)), lengths(), it + 1, refactor=False,
                         bed_pass=py_twin.reproject_bed)
    for fn in ('alignment_iteration_%d.bed', 'scaffold_length_iteration_%d', 're_counts_iteration_%d'):
        assert open(d + '/' + fn % (it + 1), 'rb').read() == get('s%d.layout.' % it + fn % (it + 1)), fn


@pytest.mark.parametrize('it', [1, 2])
def test_refactor_breaks(tmp_path, it):
    from oracle import py_twin
    from salsa_b200.host import breaks
    d = stage(tmp_path, dict(BASE, **{'scaffolds_iteration_%d.p' % it: 's%d.layout.scaffolds_iteration_%d.json' % (it, it),
                                      'misasm_iteration_%d.report' % (it + 1): 's%d.misasm_iteration_%d.report' % (it, it + 1)}))
    breaks.refactor(d, it + 1, py_twin.reproject_bed)
    for fn in ('links_avoid_iteration_%d', 'alignment_iteration_%d.bed', 'scaffold_length_iteration_%d', 're_counts_iteration_%d'):
        assert open(d + '/' + fn % (it + 1), 'rb').read() == get('s%d.refactor.' % it + fn % (it + 1)), fn
    with open('%s/scaffolds_iteration_%d.p' % (d, it), 'rb') as f:
        got = pickle.load(f)
    assert list(got.items()) == list(scaffolds('s%d.refactor.scaffolds_iteration_%d.json' % (it, it)).items())
    if it == 1:
        assert len(get('s1.refactor.links_avoid_iteration_2').splitlines()) >= 2      # the hand-flagged joins were split


def test_layout_same_scaffolds(tmp_path):
    """The greedy layout: same scaffolds as the reference's run, up to the orientation and numbering
    that the reference itself leaves to set / dict iteration order (layout_unitigs.py:595-613,621)."""
    from salsa_b200.host import scaffold_graph
    d = stage(tmp_path, dict(BASE, **{'contig_links_scaled_sorted_iteration_1': 's1.contig_links_scaled_sorted_iteration_1',
                                      'contig_links_scaled_sorted_iteration_2': 's2.contig_links_scaled_sorted_iteration_2',
                                      'scaffold_length_iteration_2': 's1.refactor.scaffold_length_iteration_2'}))
    got, _ = scaffold_graph.layout(d, 1, 1000, {})
    want = scaffolds('s1.layout.scaffolds_iteration_1.json')
    assert canonical(got) == canonical(want) and len(want) < 36
    got2, _ = scaffold_graph.layout(d, 2, 1000, scaffolds('s1.refactor.scaffolds_iteration_1.json'))
    assert canonical(got2) == canonical(scaffolds('s2.layout.scaffolds_iteration_2.json'))


def test_agp_writer(tmp_path):
    from salsa_b200.host import agp
    d = stage(tmp_path, {'assembly.cleaned.fasta': 'in.assembly.fa', 'final.p': 's2.refactor.scaffolds_iteration_2.json'})
    agp.write_outputs(d + '/assembly.cleaned.fasta', d + '/final.p', d + '/out.fasta', d + '/out.agp')
    assert open(d + '/out.agp', 'rb').read() == get('final.scaffolds_FINAL.agp')
    assert open(d + '/out.fasta', 'rb').read() == get('final.scaffolds_FINAL.fasta')


def test_cleaning_step(tmp_path):
    """correct.py (-m yes): cut contigs, rewritten FASTA and BED -- including the reference's missing final
    flush of the BED buffer (correct.py:86-91: anything after the last full million lines is dropped)."""
    import hashlib
    from salsa_b200.host import clean
    d = stage(tmp_path, {'assembly.fa': 'in.assembly.fa', 'reads.bed': 'in.reads.bed', 'input_breaks': 's0.input_breaks'})
    clean.clean(d + '/assembly.fa', d + '/input_breaks', d + '/reads.bed', d)
    assert open(d + '/alignment_iteration_1.tmp.bed', 'rb').read() == get('s0.alignment_iteration_1.tmp.bed') == b''
    assert open(d + '/asm.cleaned.fasta', 'rb').read() == get('s0.asm.cleaned.fasta')
    open(d + '/big.bed', 'wb').write(get('in.reads.bed') * 40)
    clean.clean(d + '/assembly.fa', d + '/input_breaks', d + '/big.bed', d)
    big = open(d + '/alignment_iteration_1.tmp.bed', 'rb').read()
    want = json.loads(get('s0.big40.digest'))
    assert (len(big), big.count(b'\n'), hashlib.sha256(big).hexdigest()) == (want['bytes'], want['lines'], want['sha256'])
    assert want['lines'] == 1000000
