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

import _common
from _common import engine


def sort_order(eng, node_names, u, v, r_digits, r_exp10):
    """The permutation `sort -k 5 -gr` (LC_ALL=C) puts the scaled lines in (run_pipeline.py:188,301), computed
    with two stable radix sorts on the GPU.  sort compares field 5 as a number, descending; lines whose ratios
    are equal fall to its last-resort comparison, the whole line bytewise, also reversed.  The ratio text is the
    12-digit str(float), so (decimal exponent, 12-digit integer) orders it exactly; the line bytes are decided
    by the two node names (every (u, v) occurs once, and a tab sorts below every name byte)."""
    import torch
    n = len(u)
    order = sorted(range(len(node_names)), key=lambda i: node_names[i])
    rank = np.empty(len(node_names), np.int64)
    rank[order] = np.arange(len(node_names))
    d = np.asarray(r_digits, np.int64)
    key1 = np.where(d > 0, ((np.asarray(r_exp10, np.int64) + 512) << 40) | d, 0)
    key2 = (rank[u] << 32) | rank[v]
    dev, st = eng.device, torch.cuda.current_stream().cuda_stream

    def gpu_sort(keys, vals):
        k = torch.from_numpy(~keys).to(dev)                  # descending
        ko, vo = torch.empty_like(k), torch.empty_like(vals)
        tb = C.c_size_t(0)
        eng.L.salsa_sort_pairs_u64(k.data_ptr(), vals.data_ptr(), ko.data_ptr(), vo.data_ptr(), n, None, C.byref(tb), st)
        tmp = torch.empty(tb.value + 16, dtype=torch.uint8, device=dev)
        rc = eng.L.salsa_sort_pairs_u64(k.data_ptr(), vals.data_ptr(), ko.data_ptr(), vo.data_ptr(), n, tmp.data_ptr(), C.byref(tb), st)
        if rc:
            raise RuntimeError(eng.L.salsa_last_error().decode())
        return vo

    idx = gpu_sort(key2, torch.arange(n, dtype=torch.int64, device=dev))           # ties of the next sort, in final order
    idx = gpu_sort(key1[idx.cpu().numpy()], idx)                                   # stable: ratio decides, ties stay
    return idx.cpu().numpy()


def parse_links(lines):
    """contig_links_iteration_N lines -> (node names, u, v, w, links); fast_scaled_scores.py:35-60."""
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
    return list(nodes.keys()), u, v, w, links


def parse_links_canonical(data):
    """The same result as parse_links(data.splitlines()) without a Python statement per line, for files in the
    form make_links.py writes them (exactly `u\tv\tscore\tcount\n`).  Returns None for anything else -- other
    whitespace, empty or extra fields, duplicate / self edges, negative weights -- and the caller falls back to
    the line loop, which also raises the right error."""
    if not data or any(c in data for c in (b' ', b'\r', b'\x0b', b'\x0c')) or b'\t\t' in data or b'\n\t' in data \
            or b'\t\n' in data or b'\n\n' in data or data[:1] in (b'\t', b'\n') or data[-1:] == b'\t':
        return None
    lines = data.split(b'\n')
    if lines[-1] == b'':
        lines.pop()
    if not all(ln.count(b'\t') == 3 for ln in lines):
        return None
    toks = data.split()
    n = len(lines)
    if len(toks) != 4 * n:
        return None
    nodes = {}
    try:
        u = np.array([nodes.setdefault(x, len(nodes)) for x in toks[0::4]], dtype=np.int64)
        v = np.array([nodes.setdefault(x, len(nodes)) for x in toks[1::4]], dtype=np.int64)
        w = np.array(list(map(float, toks[2::4])), dtype=np.float64)
        links = np.array(list(map(int, toks[3::4])), dtype=np.int64)
    except (ValueError, OverflowError):
        return None
    key = (np.minimum(u, v) << 32) | np.maximum(u, v)
    if (u == v).any() or (w < 0).any() or len(np.unique(key)) != n:
        return None
    return list(nodes.keys()), u.astype(np.int32), v.astype(np.int32), w, links


def scaled_text(eng, lines, want_sorted=False):
    parsed = parse_links_canonical(b''.join(lines))
    if parsed is None:
        parsed = parse_links(lines)
    return scaled_from_arrays(eng, *parsed, want_sorted=want_sorted)


def scaled_from_arrays(eng, node_names, u, v, w, links, want_sorted=False):
    """node_names: list[bytes] ('name:B' / 'name:E'), u / v: int32 node ids per link, w: float64 link weights as
    the text file carries them (12 significant digits), links: int64 counts."""
    n = len(u)
    if not n:
        return (b'', b'') if want_sorted else b''
    res = eng.scaled_scores(u, v, w, len(node_names))
    wd, wk, _ = eng.g12(w)                                         # str(w), :86
    from salsa_b200.engine import _blob
    blob, off = _blob(node_names)
    cap = n * 160 + 2 * int(np.diff(off)[u].sum() + np.diff(off)[v].sum()) + 1024
    out = np.empty(cap, dtype=np.uint8)                            # (create_string_buffer would zero it, .raw would copy all of it)
    c = lambda a, t: np.ascontiguousarray(a, t)                    # noqa: E731
    arrs = [c(u, np.int32), c(v, np.int32), c(wd, np.int64), c(wk, np.int32), c(res['b_digits'], np.int64),
            c(res['b_exp10'], np.int32), c(res['is_int1'], np.uint8), c(res['r_digits'], np.int64),
            c(res['r_exp10'], np.int32), c(links, np.int64)]
    nb = eng.L.salsa_format_scaled(blob, off.ctypes.data, *[a.ctypes.data for a in arrs], n, out.ctypes.data, cap)
    if nb < 0:
        raise RuntimeError(eng.L.salsa_last_error().decode())
    text = out[:nb].tobytes()
    if not want_sorted:
        return text
    perm = sort_order(eng, node_names, u, v, res['r_digits'], res['r_exp10'])
    arrs = [np.ascontiguousarray(a[perm]) for a in arrs]
    nb = eng.L.salsa_format_scaled(blob, off.ctypes.data, *[a.ctypes.data for a in arrs], n, out.ctypes.data, cap)
    if nb < 0:
        raise RuntimeError(eng.L.salsa_last_error().decode())
    return text, out[:nb].tobytes()


def main(argv=None):
    parser = argparse.ArgumentParser()
    parser.add_argument("-i", "--iteration", help="iteration number")
    parser.add_argument("-d", '--directory', help="output directory")
    args = parser.parse_args(argv)
    path = args.directory + '/contig_links_iteration_' + str(args.iteration)
    if not _common.KEEP_RESIDENT:
        with open(path, 'rb') as f:
            text = scaled_text(engine(), f.readlines())
    else:
        # inside run_pipeline.py: the links make_links.py just wrote are still in memory as arrays (no text to
        # re-parse), and the file the `sort -k 5 -gr` step would produce is left behind as well (GPU sort), which
        # that step's resume rule then finds in place (run_pipeline.py:186-188)
        cached = _common.cached_links(path)
        if cached is None:
            with open(path, 'rb') as f:
                data = f.read()
            cached = parse_links_canonical(data)
            if cached is None:
                parts = data.split(b'\n')                      # what readlines() yields in binary mode
                cached = parse_links([p + b'\n' for p in parts[:-1]] + ([parts[-1]] if parts[-1] else []))
        text, srt = scaled_from_arrays(engine(), *cached, want_sorted=True)
        if srt:
            with open(args.directory + '/contig_links_scaled_sorted_iteration_' + str(args.iteration), 'wb') as ofile:
                ofile.write(srt)
    with open(args.directory + '/contig_links_scaled_iteration_' + str(args.iteration), 'wb') as ofile:
        ofile.write(text)
    return 0


if __name__ == '__main__':
    sys.exit(main())
