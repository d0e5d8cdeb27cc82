#!/usr/bin/env python
"""Annotated SASS of one kernel from an ncu report: source line, share of executed warp instructions, samples.

    python tools/ncu_sass.py REPORT KERNEL_SUBSTRING > out.txt
"""
import csv
import io
import subprocess
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sass_regions as sr


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    ins = sr.load(kern)
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h = next(i for i, r in enumerate(rows) if 'Source' in r and 'Address' in r)
    hdr = rows[h]
    ci, si, src = hdr.index('Instructions Executed'), hdr.index('# Samples'), hdr.index('Source')
    body = [r for r in rows[h + 1:] if len(r) > ci]
    if len(body) != len(ins):
        print('warning: %d rows vs %d instructions' % (len(body), len(ins)))
    tot = sum(float(r[ci] or 0) for r in body)
    for r, (loc, op, ln) in zip(body, ins):
        print('%s:%d\t%7.4f%%\t%s\t%s' % (loc[0], loc[1], float(r[ci] or 0) / tot * 100, r[si], r[src].strip()))


if __name__ == '__main__':
    main()
