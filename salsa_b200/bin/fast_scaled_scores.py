#!/usr/bin/env python3
"""Drop-in for the reference's fast_scaled_scores.py (same argv, same files):

    fast_scaled_scores.py -d DIR -i N

reads DIR/contig_links_iteration_N and writes DIR/contig_links_scaled_iteration_N
(``u v w bestAlt w/bestAlt links False False``, tab separated, input order),
fast_scaled_scores.py:35-89.  Node maxima, the best-buddy alternative and the
ratios are computed on the GPU; text is assembled from the device's 12-digit
integers.
"""
import argparse
import ctypes as C
import sys

import numpy as np

from _common import engine


def scaled_text(eng, lines):
    nodes = {}
    u = np.empty(len(lines), np.int32)
    v = np.empty(len(lines), np.int32)
    w = np.empty(len(lines), np.float64)
    links = np.empty(len(lines), np.int64)
    seen = set()
    for i, ln in enumerate(lines):
        a = ln.strip().split()                                     # fast_scaled_scores.py:37
        for k in (0, 1):
            if a[k] not in nodes:
                nodes[a[k]] = len(nodes)
        u[i], v[i] = nodes[a[0]], nodes[a[1]]
        w[i] = float(a[2])                                         # :48
        links[i] = int(a[3])                                       # :59
        e = (min(u[i], v[i]), max(u[i], v[i]))
        if e in seen or u[i] == v[i]:
            raise ValueError('duplicate or self edge in contig_links: %r (the link graph must be simple)' % ln)
        seen.add(e)
        if w[i] < 0:
            raise ValueError('negative link weight: %r' % ln)
    if not len(lines):
        return b''
    nn = len(nodes)
    res = eng.scaled_scores(u, v, w, nn)
    wd, wk, _ = eng.g12(w)                                         # str(w), :86
    from salsa_b200.engine import _blob
    blob, off = _blob(list(nodes.keys()))
    cap = len(lines) * 160 + 2 * sum(len(x) for x in lines) + 1024
    out = C.create_string_buffer(cap)
    c = lambda a, t: np.ascontiguousarray(a, t)                    # noqa: E731
    arrs = [c(u, np.int32), c(v, np.int32), c(wd, np.int64), c(wk, np.int32), c(res['b_digits'], np.int64),
            c(res['b_exp10'], np.int32), c(res['is_int1'], np.uint8), c(res['r_digits'], np.int64),
            c(res['r_exp10'], np.int32), c(links, np.int64)]
    nb = eng.L.salsa_format_scaled(blob, off.ctypes.data, *[a.ctypes.data for a in arrs], len(lines), out, cap)
    if nb < 0:
        raise RuntimeError(eng.L.salsa_last_error().decode())
    return out.raw[:nb]


def main(argv=None):
    parser = argparse.ArgumentParser()
    parser.add_argument("-i", "--iteration", help="iteration number")
    parser.add_argument("-d", '--directory', help="output directory")
    args = parser.parse_args(argv)
    with open(args.directory + '/contig_links_iteration_' + str(args.iteration), 'rb') as f:
        lines = f.readlines()
    text = scaled_text(engine(), lines)
    with open(args.directory + '/contig_links_scaled_iteration_' + str(args.iteration), 'wb') as ofile:
        ofile.write(text)
    return 0


if __name__ == '__main__':
    sys.exit(main())
