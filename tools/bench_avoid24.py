import sys, time, json
sys.path.insert(0, '.')
import numpy as np, torch
from salsa_b200 import synth
from salsa_b200.engine import Engine
dev = torch.device('cuda', 0)
eng = Engine(0)
asm = synth.make_assembly(ncontig=24, genome=3_000_000_000, len_lo=40_000_000, len_hi=240_000_000, nchrom=24, seed=3)
pairs = 100_000_000
buf = torch.empty(int(pairs * 2.02 * 64) + (1 << 20), dtype=torch.uint8, device=dev)
pos = done = 0
while done < pairs:
    m = min(20_000_000, pairs - done)
    rec = synth.gen_records(asm, m, seed=100 + done, device=dev, read_id_base=done, noise=0.2)
    out = synth.format_bed(rec, out=buf[pos:]); pos += out.numel(); done += m; del rec, out
bed = buf[:pos]
n = asm.ncontig
one = np.ones(n, np.int64)
for label, avoid in (('none', []), ('12 disjoint pairs over all 24 scaffolds', [(asm.names[2 * i], asm.names[2 * i + 1]) for i in range(12)]),
                     ('3 pairs', [(asm.names[2 * i], asm.names[2 * i + 1]) for i in range(3)])):
    eng.set_contigs(asm.names, asm.lengths, re_left=one, re_right=one, avoid=avoid)
    eng.begin_links('dense')
    eng.ingest_device(bed); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); eng.ingest_device(bed); b.record(); torch.cuda.synchronize()
    st = eng.stats_read()
    print(json.dumps({'avoid': label, 'ms': a.elapsed_time(b), 'bytes': int(bed.numel()), 'counted': int(st.pairs)}), flush=True)
