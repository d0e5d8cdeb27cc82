#!/usr/bin/env python
"""Static SASS instruction counts of one kernel, per source line and per named source region
(no GPU needed: cuobjdump + nvdisasm on the built library, -lineinfo).

    python tools/sass_regions.py KERNEL_SUBSTRING [file:lo-hi ...] [--ops] [--lines N]
"""
import collections
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load(kern, so=None):
    so = so or os.path.join(ROOT, 'salsa_b200', 'libsalsa_b200.so')
    d = tempfile.mkdtemp()
    subprocess.run(['cuobjdump', '-xelf', 'all', so], cwd=d, check=True, capture_output=True)
    out = []
    for f in sorted(os.listdir(d)):
        if not f.endswith('.cubin'):
            continue
        txt = subprocess.run(['nvdisasm', '-g', '-c', os.path.join(d, f)], capture_output=True, text=True).stdout
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
            m = re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)', ln)
            if m:
                out.append((cur, m.group(2), ln.strip()))
        if out:
            break
    return out


def main():
    kern = sys.argv[1]
    regions = [a for a in sys.argv[2:] if re.match(r'.+:\d+-\d+$', a)]
    ins = load(kern)
    print('kernel %s: %d SASS instructions' % (kern, len(ins)))
    per = collections.Counter(loc for loc, _, _ in ins)
    for r in regions:
        f, ab = r.rsplit(':', 1)
        a, b = map(int, ab.split('-'))
        sel = [(loc, op) for loc, op, _ in ins if loc[0] == f and a <= loc[1] <= b]
        print('%-28s %5d' % (r, len(sel)), end='')
        if '--ops' in sys.argv:
            c = collections.Counter(op.split('.')[0] for _, op in sel)
            print('  ' + ' '.join('%s:%d' % kv for kv in c.most_common(12)), end='')
        print()
    if '--lines' in sys.argv:
        n = int(sys.argv[sys.argv.index('--lines') + 1])
        for loc, v in per.most_common(n):
            print('%5d  %s:%d' % (v, loc[0], loc[1]))


if __name__ == '__main__':
    main()
