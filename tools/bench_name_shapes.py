#!/usr/bin/env python
"""Ingest throughput for input shapes off the generator's default: long contig names (fast-path windows are
re-centred past 16 bytes), short read names.  20 M pairs, one B200; prints one line per shape."""
import sys, time
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from salsa_b200 import synth
from salsa_b200.engine import Engine
dev = torch.device('cuda', 0)
eng = Engine(0)
asm = synth.make_assembly(seed=5, ncontig=3000, genome=3_100_000_000, len_lo=50_000, len_hi=20_000_000, nchrom=23)
rec = synth.gen_records(asm, 20_000_000, seed=3, device=dev)
bed = synth.format_bed(rec)
base = bytes(bed.cpu().numpy())
del bed
def run(tag, text, names):
    t = torch.frombuffer(bytearray(text), dtype=torch.uint8).to(dev)
    one = np.ones(len(names), np.int64)
    eng.set_contigs(names, asm.lengths, one, one)
    eng.begin_links('dense')
    for _ in range(2): eng.table.zero_(); eng.ingest_device(t)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(5): eng.table.zero_(); eng.ingest_device(t)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 5
    st = eng.stats_read()
    print('%-28s %6.2f ms  %7.1f GB/s  lines %d pairs %d' % (tag, dt * 1e3, len(text) / dt / 1e9, st.lines // 7, st.pairs // 7), flush=True)
run('ctg%06d (9 B names)', base, asm.names)
long_names = [b'NODE_%06d_length_1234567_cov_12.345' % i for i in range(3000)]
run('35 B names', base.replace(b'ctg', b'NODE_').replace(b'\t', b'_length_1234567_cov_12.345\t', 0) if False else b'\n'.join(l.replace(b'ctg', b'NODE_', 1).replace(b'\t', b'_length_1234567_cov_12.345\t', 1) for l in base.split(b'\n')), long_names)
mid_names = [b'tig%08d_pilon_x' % i for i in range(3000)]
run('19 B names', b'\n'.join(l.replace(b'ctg', b'tig00', 1).replace(b'\t', b'_pilon_x\t', 1) for l in base.split(b'\n')), mid_names)
run('short reads r123/1', base.replace(b'read0000', b'r'), asm.names)
run('38 B read names', base.replace(b'read0000', b'A00123:45:HXXXXXXXX:1:1101:12345:'), asm.names)
run('4-column BED', b'\n'.join(l.rsplit(b'\t', 2)[0] for l in base.split(b'\n') if l) + b'\n', asm.names)
