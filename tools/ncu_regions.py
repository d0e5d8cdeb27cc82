#!/usr/bin/env python
"""Per-region instruction shares of one kernel in an ncu report (tools/ncu_lines.py grouped by source ranges).

    python tools/ncu_regions.py REPORT KERNEL LINES name=file:lo-hi ...
LINES = number of BED lines the profiled launch scanned (for the instructions-per-line column).
"""
import collections
import re
import subprocess
import sys
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    rep, kern, nlines = sys.argv[1], sys.argv[2], float(sys.argv[3])
    reg = collections.OrderedDict()
    for a in sys.argv[4:]:
        nm, r = a.split('=')
        f, ab = r.rsplit(':', 1)
        lo, hi = map(int, ab.split('-'))
        reg[nm] = (f, lo, hi)
    out = subprocess.run([sys.executable, os.path.join(ROOT, 'tools', 'ncu_lines.py'), rep, kern, '--top', '100000'],
                         capture_output=True, text=True).stdout
    tot = float(re.search(r'total warp instructions ([\d.e+]+)', out).group(1))
    acc, other = collections.Counter(), collections.Counter()
    for ln in out.splitlines():
        m = re.match(r'\s*([\d.]+)% inst\s+([\d.]+)% samples\s+thr/inst\s+([\d.]+)\s+(\S+):(\d+)', ln)
        if not m:
            continue
        pct, f, l = float(m.group(1)), m.group(4), int(m.group(5))
        for k, (ff, a, b) in reg.items():
            if f == ff and a <= l <= b:
                acc[k] += pct
                break
        else:
            other[f] += pct
    print('total %.3e warp instructions, %.2f per line' % (tot, tot / nlines))
    for k in reg:
        print('%-28s %5.1f%%  %5.2f instr/line' % (k, acc[k], acc[k] / 100 * tot / nlines))
    for k, v in other.most_common(8):
        print('other %-22s %5.1f%%  %5.2f' % (k, v, v / 100 * tot / nlines))


if __name__ == '__main__':
    main()
