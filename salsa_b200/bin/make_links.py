#!/usr/bin/env python3
"""Drop-in for the reference's make_links.py (same argv, same files):

    make_links.py -b BED -d DIR -i N [-x DUP]

reads DIR/scaffold_length_iteration_N, DIR/re_counts_iteration_N and (N > 1)
DIR/links_avoid_iteration_N, writes DIR/contig_links_iteration_N
(``u:X\\tv:Y\\tscore\\tcount``) and prints the number of links, like
make_links.py:159.  The BED scan, pairing, histogram and scoring run on the GPU
(libsalsa_b200); lines are written in (first, second, ends) order -- the
reference's own order is Python-2 dict order (make_links.py:161).
``-x`` is accepted and ignored: in the reference it is dead code
(make_links.py:20-22 compares a str with an int).
Any condition that makes the reference raise exits non-zero without writing
the output file (run_pipeline.py:166-170 treats that as a failed step).
"""
import argparse
import sys

import numpy as np

import _common
from _common import engine


def load_tables(directory, iteration):
    from salsa_b200.engine import length_floor
    names, lens = {}, []
    with open(directory + '/scaffold_length_iteration_' + iteration, 'rb') as f:      # make_links.py:37-41
        for line in f:
            a = line.split()
            names[a[0]] = length_floor(a[1].decode('latin-1'))
    nm = list(names.keys())
    lens = np.array([names[k] for k in nm], dtype=np.int64)
    idx = {k: i for i, k in enumerate(nm)}
    has_re = np.zeros(len(nm), dtype=np.uint8)
    rl = np.zeros(len(nm), dtype=np.int64)
    rr = np.zeros(len(nm), dtype=np.int64)
    with open(directory + '/re_counts_iteration_' + iteration, 'rb') as f:            # make_links.py:43-48
        for line in f:
            a = line.split()
            left, right = int(a[1]), int(a[2])
            if left < 0 or right < 0:
                raise ValueError('negative RE count for %r' % a[0])
            if a[0] in idx:
                i = idx[a[0]]
                has_re[i], rl[i], rr[i] = 1, left, right
    avoid = []
    if int(iteration) > 1:                                                            # make_links.py:29-35
        with open(directory + '/links_avoid_iteration_' + iteration, 'rb') as f:
            for line in f:
                a = line.split()
                avoid.append((a[0], a[1]))
    return nm, lens, rl, rr, has_re, avoid


def main(argv=None):
    parser = argparse.ArgumentParser()
    parser.add_argument("-b", "--mapping", help="mapping of read to contigs in bed format")
    parser.add_argument("-i", "--iteration", help="iteration number")
    parser.add_argument("-d", '--directory', help="output directory")
    parser.add_argument("-x", "--dup", help="Duplicated links")
    args = parser.parse_args(argv)
    iteration = str(args.iteration)
    nm, lens, rl, rr, has_re, avoid = load_tables(args.directory, iteration)
    eng = engine()
    eng.set_contigs(nm, lens, rl, rr, has_re, avoid)
    print('bedfile started', file=sys.stderr)
    from salsa_b200.reproject import resident_bed
    try:
        # left in HBM by the re-projection that wrote it, or loaded now and kept for the re-projections to come
        bed = resident_bed(args.mapping, eng.device, load=_common.KEEP_RESIDENT)
    except MemoryError:
        bed = None
    # dense table, or (fragmented assemblies) a sparse one that grows like the reference's dict; stats_read inside
    # raises DataError where the reference would raise
    if bed is not None:
        eng.ingest_links(lambda: eng.ingest_device(bed))
    else:
        eng.ingest_links(lambda: eng.ingest_file(args.mapping))   # streamed through a staging buffer; nothing stays on the device
    print('bedfile loaded', file=sys.stderr)
    sc = eng.score_links()
    print(sc['n'])
    out_path = args.directory + '/contig_links_iteration_' + iteration
    with open(out_path, 'wb') as ofile:
        ofile.write(eng.link_text(sc))
    if _common.KEEP_RESIDENT and sc['n']:
        # the next step re-reads this file; hand it the same numbers as arrays (w = float(str(score)), the value
        # the text carries; node = contig end)
        nodes = [nm + b':' + e for nm in eng.names for e in (b'B', b'E')]
        u = (sc['first'] * 2 + (sc['ends'] >> 1)).astype(np.int32)
        v = (sc['second'] * 2 + (sc['ends'] & 1)).astype(np.int32)
        w = np.where(sc['is_int0'] != 0, 0.0, sc['w12'].cpu().numpy())
        _common.remember_links(out_path, nodes, u, v, w, np.asarray(sc['count'], np.int64))
    return 0


if __name__ == '__main__':
    sys.exit(main())
