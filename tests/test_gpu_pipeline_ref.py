"""GPU side of tests/test_pipeline_ref.py: the CUDA steps fed the REFERENCE's inputs must write the
REFERENCE's outputs (tests/golden/pipeline_ref/, produced by the reference's own programs)."""
import io
import os
import pickle
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import test_pipeline_ref as T  # noqa: E402

BIN = os.path.join(T.ROOT, 'salsa_b200', 'bin')


@pytest.fixture(scope='module')
def eng():
    sys.path.insert(0, BIN)
    from _common import engine
    return engine()


def test_hot_steps(eng, tmp_path):
    import RE_sites
    import fast_scaled_scores
    import make_links
    d = T.stage(tmp_path, dict(T.BASE, **{'a.fa': 'in.assembly.fa',
                                          'scaffold_length_iteration_2': 's1.refactor.scaffold_length_iteration_2',
                                          're_counts_iteration_2': 's1.refactor.re_counts_iteration_2',
                                          'links_avoid_iteration_2': 's1.refactor.links_avoid_iteration_2',
                                          'alignment_iteration_2.bed': 's1.refactor.alignment_iteration_2.bed'}))
    buf = io.StringIO()
    assert RE_sites.main(['-a', d + '/a.fa', '-e', 'GATC'], out=buf) == 0
    assert buf.getvalue().encode() == T.get('s1.re_counts_iteration_1')
    for it in (1, 2):
        assert make_links.main(['-b', '%s/alignment_iteration_%d.bed' % (d, it), '-d', d, '-i', str(it)]) == 0
        got = open('%s/contig_links_iteration_%d' % (d, it), 'rb').read().splitlines(keepends=True)
        assert sorted(got) == sorted(T.get('s%d.contig_links_iteration_%d' % (it, it)).splitlines(keepends=True))
        open('%s/contig_links_iteration_%d' % (d, it), 'wb').write(T.get('s%d.contig_links_iteration_%d' % (it, it)))
        assert fast_scaled_scores.main(['-d', d, '-i', str(it)]) == 0
        assert open('%s/contig_links_scaled_iteration_%d' % (d, it), 'rb').read() == T.get('s%d.contig_links_scaled_iteration_%d' % (it, it))
        r = subprocess.run(['sort', '-k', '5', '-gr', '%s/contig_links_scaled_iteration_%d' % (d, it)], capture_output=True,
                           env=dict(os.environ, LC_ALL='C'))
        assert r.stdout == T.get('s%d.contig_links_scaled_sorted_iteration_%d' % (it, it))


@pytest.mark.parametrize('it', [1, 2])
def test_update_bed_after_layout(eng, tmp_path, it):
    from salsa_b200 import reproject as gpu_reproject
    from salsa_b200.host import reproject
    d = T.stage(tmp_path, T.BASE)
    reproject.update_bed(d, T.scaffolds('s%d.layout.scaffolds_iteration_%d.json' % (it, it)), T.lengths(), it + 1, refactor=False,
                         bed_pass=gpu_reproject.bed_pass(eng))
    for fn in ('alignment_iteration_%d.bed', 'scaffold_length_iteration_%d', 're_counts_iteration_%d'):
        assert open(d + '/' + fn % (it + 1), 'rb').read() == T.get('s%d.layout.' % it + fn % (it + 1)), fn


@pytest.mark.parametrize('it', [1, 2])
def test_refactor_breaks(eng, tmp_path, it):
    from salsa_b200 import reproject as gpu_reproject
    from salsa_b200.host import breaks
    d = T.stage(tmp_path, dict(T.BASE, **{'scaffolds_iteration_%d.p' % it: 's%d.layout.scaffolds_iteration_%d.json' % (it, it),
                                          'misasm_iteration_%d.report' % (it + 1): 's%d.misasm_iteration_%d.report' % (it, it + 1)}))
    breaks.refactor(d, it + 1, gpu_reproject.bed_pass(eng))
    for fn in ('links_avoid_iteration_%d', 'alignment_iteration_%d.bed', 'scaffold_length_iteration_%d', 're_counts_iteration_%d'):
        assert open(d + '/' + fn % (it + 1), 'rb').read() == T.get('s%d.refactor.' % it + fn % (it + 1)), fn
    with open('%s/scaffolds_iteration_%d.p' % (d, it), 'rb') as f:
        assert list(pickle.load(f).items()) == list(T.scaffolds('s%d.refactor.scaffolds_iteration_%d.json' % (it, it)).items())
