"""End-to-end parity on the GPU box: the drop-in pipeline with the CUDA steps against the same
pipeline with the oracle / compiled-reference steps (same host-side code in both runs).  Every
intermediate file must agree (order-insensitively only where the reference's own order is a hash
order), contig_links_scaled_sorted_iteration_N and scaffolds_FINAL.agp byte for byte."""
import os
import subprocess
import sys

import pytest

import pipeline_util as pu

pytestmark = pytest.mark.gpu


def run(bin_dir, data, out, extra):
    cmd = [sys.executable, os.path.join(pu.BIN, 'run_pipeline.py'), '--bin', bin_dir, '-a', data + '/assembly.fa', '-l',
           data + '/contigs.len', '-b', data + '/reads.bed', '-o', out] + extra
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=3000)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]


def compare(out_g, out_o):
    g, o = pu.comparable(out_g), pu.comparable(out_o)
    assert sorted(g) == sorted(o)
    for fn in g:
        assert g[fn] == o[fn], fn
    assert 'scaffolds_FINAL.agp' in g and len(g['scaffolds_FINAL.agp']) > 100
    return g


def test_config1_like(tmp_path, oracle_lib):
    """BASELINE configs[0] shape, scaled down: small contigs (no coverage work triggers), -e GATC -i 3."""
    data = str(tmp_path / 'data')
    pu.write_dataset(data, ncontig=200, genome=2_000_000, len_lo=2_000, len_hi=40_000, npairs=200_000, seed=20261107, nchrom=4)
    obin = pu.make_oracle_bin(str(tmp_path / 'obin'))
    args = ['-e', 'GATC', '-i', '3', '-m', 'no']
    run(pu.BIN, data, str(tmp_path / 'gpu'), args)
    run(obin, data, str(tmp_path / 'ora'), args)
    g = compare(str(tmp_path / 'gpu'), str(tmp_path / 'ora'))
    assert any(fn.startswith('links_avoid_iteration_') for fn in g)


def test_spawned_steps_equal_in_process_steps(tmp_path):
    """--subprocess: every step its own process, as the reference runs them (files are the only hand-off: no
    resident BEDs, no write-behind, coreutils sort) -- same directory content as the in-process default."""
    data = str(tmp_path / 'data')
    pu.write_dataset(data, ncontig=60, genome=700_000, len_lo=2_000, len_hi=40_000, npairs=60_000, seed=5, nchrom=3)
    args = ['-e', 'GATC', '-i', '2', '-m', 'no']
    run(pu.BIN, data, str(tmp_path / 'inproc'), args)
    run(pu.BIN, data, str(tmp_path / 'spawned'), args + ['--subprocess'])
    compare(str(tmp_path / 'inproc'), str(tmp_path / 'spawned'))
    assert not [f for f in os.listdir(str(tmp_path / 'inproc')) if f.endswith('.part')]


def test_dnase_with_misassembly_correction(tmp_path, oracle_lib):
    """BASELINE configs[2] shape, scaled down: -e DNASE (RE counts = len / 2, RE_sites.py:39-43) with -m yes on a genome
    whose contigs are long enough for both coverage programs to do real work, 3 iterations."""
    data = str(tmp_path / 'data')
    pu.write_dataset(data, ncontig=40, genome=60_000_000, len_lo=400_000, len_hi=3_500_000, npairs=2_000_000, seed=77, nchrom=3)
    obin = pu.make_oracle_bin(str(tmp_path / 'obin'))
    args = ['-e', 'DNASE', '-i', '3', '-m', 'yes']
    run(pu.BIN, data, str(tmp_path / 'gpu'), args)
    run(obin, data, str(tmp_path / 'ora'), args)
    g = compare(str(tmp_path / 'gpu'), str(tmp_path / 'ora'))
    assert 'input_breaks' in g and 'misasm_iteration_2.report' in g and 'scaffolds_FINAL.fasta' in g


def test_misassembly_and_gfa(tmp_path, oracle_lib):
    """>= 1 Mb contigs so break_contigs_start / break_contigs do real work; -m yes, -g GFA, Arima enzymes."""
    data = str(tmp_path / 'data')
    pu.write_dataset(data, ncontig=14, genome=24_000_000, len_lo=900_000, len_hi=4_000_000, npairs=1_100_000, seed=31, nchrom=2,
                     with_gfa=True)
    obin = pu.make_oracle_bin(str(tmp_path / 'obin'))
    args = ['-e', 'GATC,GANTC', '-i', '2', '-m', 'yes', '-g', data + '/assembly.gfa']
    run(pu.BIN, data, str(tmp_path / 'gpu'), args)
    run(obin, data, str(tmp_path / 'ora'), args)
    g = compare(str(tmp_path / 'gpu'), str(tmp_path / 'ora'))
    assert 'input_breaks' in g and 'misasm_iteration_2.report' in g
