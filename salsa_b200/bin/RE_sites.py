#!/usr/bin/env python3
"""Drop-in for the reference's RE_sites.py (same argv, same stdout):

    RE_sites.py -a FASTA -e ENZYME[,ENZYME...] | DNASE   > re_counts_iteration_1

prints ``id left right`` per record (RE_sites.py:88; space separated), in
first-appearance order (the reference prints in Python-2 dict order,
RE_sites.py:61, which no consumer depends on).  Exit status 1 for a ':' in a
header (RE_sites.py:13-15) or a malformed enzyme string (RE_sites.py:48-50).
"""
import argparse
import os
import sys

import numpy as np

from _common import engine


def main(argv=None, out=None):
    parser = argparse.ArgumentParser()
    parser.add_argument("-a", "--assembly", help="assembled contigs")
    parser.add_argument("-e", "--enzyme", help="restriction enzyme")
    args = parser.parse_args(argv)
    from salsa_b200.re_sites import FastaError, count_sites
    out = out or sys.stdout
    size = os.path.getsize(args.assembly)
    fasta = np.memmap(args.assembly, dtype=np.uint8, mode='r') if size else np.zeros(0, np.uint8)
    eng = engine()
    dev_copy = None
    if size > (64 << 20):
        # a 3 Gb assembly: four readers side by side into pinned buffers (one thread copies out of the page cache at a
        # tenth of what the host link takes); the header lines are read from the mapping
        from salsa_b200.reproject import _stream_in
        dev_copy = _stream_in(os.path.realpath(args.assembly), size, eng.device)
    try:
        names, left, right, seqlen = count_sites(eng, fasta, args.enzyme, device_copy=dev_copy)
    except FastaError as e:
        print(str(e), file=sys.stderr)
        return 1
    out.write(''.join('%s %d %d\n' % (nm.decode('latin-1'), l, r) for nm, l, r in zip(names, left, right)))
    return 0


if __name__ == '__main__':
    sys.exit(main())
