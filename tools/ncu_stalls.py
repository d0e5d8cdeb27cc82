#!/usr/bin/env python
"""Stall reasons of one kernel in an ncu report, summed over its SASS, and the instructions that collect the most
stall samples (with the source line from nvdisasm -g when the library is the one that was profiled).

    python tools/ncu_stalls.py REPORT [--top N]
"""
import collections
import csv
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[sys.argv.index('--top') + 1]) if '--top' in sys.argv else 30
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, data = rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
    tot = collections.Counter()
    for r in data:
        for s in stalls:
            try:
                tot[s] += int(r[ix[s]])
            except (ValueError, IndexError):
                pass
    S = sum(tot.values())
    print('total samples', S)
    for s, v in tot.most_common(10):
        print('%-26s %6.2f%%' % (s, 100 * v / S))
    n = lambda r, k: int(r[ix[k]] or 0)          # noqa: E731
    for pos, r in sorted(enumerate(data), key=lambda t: -n(t[1], '# Samples'))[:top]:
        d = max(stalls, key=lambda s: n(r, s))
        prev = data[pos - 1][ix['Source']].strip()[:44] if pos else ''
        print('%6d %5.2f%%  %-58s %-22s exec=%-9s | after: %s' % (n(r, '# Samples'), 100 * n(r, '# Samples') / S, r[ix['Source']].strip()[:58], d,
                                                                 r[ix['Instructions Executed']], prev))


if __name__ == '__main__':
    main()
