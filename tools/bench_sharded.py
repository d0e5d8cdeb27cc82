#!/usr/bin/env python
"""Wall time of the two paths that shard by contig, on BASELINE configs[1]'s assembly
(3.1 Gb, 3,000 contigs): RE-site counting (dist.count_sites_sharded) and break_contigs_start
(dist.coverage_sharded).  Run under torchrun with N ranks (or plainly for N = 1); rank 0 prints one JSON line.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/bench_sharded.py [pairs]
"""
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as td  # noqa: E402


def main():
    pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 50_000_000
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    os.environ['SALSA_B200_DEVICE'] = str(local)
    if world > 1:
        td.init_process_group('nccl', device_id=torch.device('cuda', local))
    from salsa_b200 import dist, synth, _lib
    from salsa_b200.engine import Engine
    from salsa_b200.re_sites import count_sites
    eng = Engine(local)
    cfg = dict(synth.CONFIGS['cfg2'])
    asm = synth.make_assembly(seed=20261108, **cfg)
    fasta = np.frombuffer(synth.gen_fasta(asm), dtype=np.uint8)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            td.barrier()

    out = {'world': world, 'fasta_bytes': int(fasta.size)}
    # ---- RE sites: each rank counts the records of its byte range (host memory -> its GPU -> kernels), all-gather
    fn = lambda shard, enz: count_sites(eng, np.ascontiguousarray(shard), enz)   # noqa: E731
    dist.count_sites_sharded(fn, fasta, 'GATC,GANTC')
    barrier()
    t0 = time.time()
    names, left, right, seqlen = dist.count_sites_sharded(fn, fasta, 'GATC,GANTC')
    barrier()
    out['re_sites_s'] = round(time.time() - t0, 3)
    out['re_contigs'] = len(names)
    out['re_checksum'] = int(np.sum(left) + 3 * np.sum(right))
    # ---- coverage: every rank holds the BED, scans it, and analyses ITS contigs
    d = tempfile.mkdtemp()
    lens_path = os.path.join(d, 'contigs.len')
    if rank == 0:
        open(lens_path + '.tmp', 'wb').write(asm.lengths_text())
        os.replace(lens_path + '.tmp', lens_path)
    else:
        open(lens_path, 'wb').write(asm.lengths_text())
    dev = torch.device('cuda', local)
    buf = torch.empty(int(pairs * 2.02 * 48) + (1 << 20), dtype=torch.uint8, device=dev)
    pos = done = 0
    while done < pairs:
        m = min(20_000_000, pairs - done)
        rec = synth.gen_records(asm, m, seed=900 + done, device=dev, read_id_base=done)
        o = synth.format_bed(rec, out=buf[pos:])
        pos += o.numel()
        done += m
        del rec, o
    bed = buf[:pos]
    import ctypes as C
    L = _lib.lib()

    def run_fn(lp):
        outp = os.path.join(d, 'report.rank%d' % rank)
        argv = ['-a', 'resident.bed', '-l', lp, '-s', '100']
        arr = (C.c_char_p * len(argv))(*[a.encode() for a in argv])
        rc = L.salsa_break_contigs_run(0, len(argv), arr, bed.data_ptr(), bed.numel(), outp.encode())
        assert rc == 0, L.salsa_last_error()
        return open(outp, 'rb').read()

    dist.coverage_sharded(run_fn, lens_path, False, d)
    barrier()
    t0 = time.time()
    rep = dist.coverage_sharded(run_fn, lens_path, False, d)
    barrier()
    out['break_contigs_start_s'] = round(time.time() - t0, 3)
    out['bed_bytes'] = int(bed.numel())
    out['report_lines'] = rep.count(b'\n')
    out['report_checksum'] = int(np.frombuffer(rep, dtype=np.uint8).astype(np.int64).sum())
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        td.destroy_process_group()


if __name__ == '__main__':
    main()
