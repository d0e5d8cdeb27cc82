"""CPU end-to-end run of the drop-in pipeline with the ORACLE as the hot path (no GPU needed):
exercises the host-side steps (layout, re-projection, misjoin splitting, AGP) and the driver's
iteration / resume logic."""
import os
import subprocess
import sys

import pipeline_util as pu


def run(bin_dir, data, out, extra):
    cmd = [sys.executable, os.path.join(pu.BIN, 'run_pipeline.py'), '--bin', bin_dir, '-a', data + '/assembly.fa', '-l',
           data + '/contigs.len', '-b', data + '/reads.bed', '-o', out] + extra
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=1200)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    return r


def test_pipeline_with_oracle_steps(tmp_path, oracle_lib):
    data = str(tmp_path / 'data')
    asm = pu.write_dataset(data, ncontig=60, genome=1_200_000, len_lo=6_000, len_hi=60_000, npairs=120_000, seed=5)
    obin = pu.make_oracle_bin(str(tmp_path / 'obin'))
    out = str(tmp_path / 'out')
    run(obin, data, out, ['-e', 'GATC', '-i', '3', '-m', 'no', '-c', '1000'])
    agp = open(out + '/scaffolds_FINAL.agp').read().splitlines()
    placed = [ln.split('\t')[5] for ln in agp if ln.split('\t')[4] == 'W']
    assert sorted(placed) == sorted(n.decode() for n in asm.names)          # every contig exactly once
    nscaf = len({ln.split('\t')[0] for ln in agp})
    assert nscaf < asm.ncontig / 2                                          # Hi-C links did join contigs
    assert os.path.getsize(out + '/scaffolds_FINAL.fasta') > asm.genome
    for it in (1, 2):
        assert os.path.isfile(out + '/contig_links_scaled_sorted_iteration_%d' % it)
        assert os.path.isfile(out + '/misasm_iteration_%d.report' % (it + 1))
    # resume: a second run over the same directory redoes nothing but the final write and gives the same AGP
    before = open(out + '/scaffolds_FINAL.agp').read()
    os.remove(out + '/scaffolds_FINAL.agp')
    run(obin, data, out, ['-e', 'GATC', '-i', '3', '-m', 'no', '-c', '1000'])
    assert open(out + '/scaffolds_FINAL.agp').read() == before
