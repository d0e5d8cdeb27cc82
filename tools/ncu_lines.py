#!/usr/bin/env python
"""Join an ncu report's SASS page with nvdisasm line info -> per-source-line
instruction and stall-sample shares (ncu's CSV source page carries no metrics
in its CUDA view).

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep bed_ingest_kernelILi0 [--top 40]
"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    top = int(sys.argv[sys.argv.index('--top') + 1]) if '--top' in sys.argv else 40
    so = os.path.join(ROOT, 'salsa_b200', 'libsalsa_b200.so')
    d = tempfile.mkdtemp()
    subprocess.run(['cuobjdump', '-xelf', 'all', so], cwd=d, check=True, capture_output=True)
    lines_of = []          # source line of the i-th instruction of the kernel
    for f in sorted(os.listdir(d)):
        if not f.endswith('.cubin'):
            continue
        txt = subprocess.run(['nvdisasm', '-g', '-c', os.path.join(d, f)], capture_output=True, text=True).stdout
        if kern not in txt:
            continue
        inside, cur = False, (None, 0)
        for ln in txt.splitlines():
            if ln.startswith('//---') and '.text.' in ln:
                inside = kern in ln
                continue
            if not inside:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
            if m:
                cur = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            if re.match(r'\s+/\*[0-9a-f]{4,}\*/', ln):
                lines_of.append((cur, ln.strip()))
        if lines_of:
            break
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h = next(i for i, r in enumerate(rows) if 'Source' in r and 'Address' in r)
    hdr = rows[h]
    ci, ti, si = hdr.index('Instructions Executed'), hdr.index('Thread Instructions Executed'), hdr.index('# Samples')
    body = [r for r in rows[h + 1:] if len(r) > ci]
    if len(body) != len(lines_of):
        print('warning: %d SASS rows in report vs %d in nvdisasm' % (len(body), len(lines_of)))
    agg = defaultdict(lambda: [0.0, 0.0, 0.0])
    for r, (loc, _) in zip(body, lines_of):
        a = agg[loc]
        a[0] += float(r[ci] or 0); a[1] += float(r[ti] or 0); a[2] += float(r[si] or 0)
    tot = sum(a[0] for a in agg.values()); tots = sum(a[2] for a in agg.values())
    src = {}
    print('total warp instructions %.3e, samples %d' % (tot, tots))
    for loc, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        fn, ln = loc
        if fn and fn not in src:
            p = os.path.join(ROOT, 'salsa_b200', 'csrc', fn)
            src[fn] = open(p).read().splitlines() if os.path.exists(p) else []
        text = src.get(fn, [])[ln - 1].strip() if fn and 0 < ln <= len(src.get(fn, [])) else ''
        print('%5.1f%% inst  %5.1f%% samples  thr/inst %4.1f  %s:%d  %s' % (a[0] / tot * 100, a[2] / max(tots, 1) * 100,
                                                                          a[1] / max(a[0], 1), fn, ln, text[:90]))


if __name__ == '__main__':
    main()
