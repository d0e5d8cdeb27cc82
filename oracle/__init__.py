"""oracle -- ctypes face of oracle/salsa_oracle.c (the CPU restatement).

TEST INFRASTRUCTURE ONLY.  Importers allowed: tests/, __graft_entry__.smoke(),
bench.py's cpu_baseline / --impl reference legs.  The product package
(salsa_b200/) never imports this module.

Parity status is stated in oracle/salsa_oracle.c's header: the Python steps
are pinned against vectors produced by the reference's own scripts under
tests/golden/ref_py2emu.py (the reference has no tests of its own), the
coverage steps against the compiled reference in oracle/_ref/.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

REF_DIR = os.path.join(_HERE, '_ref')


def build(quiet=True):
    """Compile liboracle.so and, when /root/reference exists, oracle/_ref."""
    r = subprocess.run(['make', '-C', _HERE], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError('oracle build failed:\n' + r.stdout + r.stderr)
    if not quiet:
        print(r.stdout)


class _Links(C.Structure):
    _fields_ = [('n', C.c_int64), ('first', C.POINTER(C.c_int32)), ('second', C.POINTER(C.c_int32)),
                ('ends', C.POINTER(C.c_uint8)), ('count', C.POINTER(C.c_int64)),
                ('norm', C.POINTER(C.c_double)), ('score', C.POINTER(C.c_double)),
                ('score_is_int0', C.POINTER(C.c_uint8)), ('nlines', C.c_int64), ('npairs_counted', C.c_int64)]


class _Re(C.Structure):
    _fields_ = [('n', C.c_int64), ('names_blob', C.POINTER(C.c_char)), ('name_off', C.POINTER(C.c_int64)),
                ('left', C.POINTER(C.c_int64)), ('right', C.POINTER(C.c_int64)), ('seqlen', C.POINTER(C.c_int64))]


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, 'liboracle.so')
        if not os.path.exists(path):
            build()
        _LIB = C.CDLL(path)
        _LIB.oracle_last_error.restype = C.c_char_p
        _LIB.oracle_round12.restype = C.c_double
        _LIB.oracle_round12.argtypes = [C.c_double]
    return _LIB


class OracleError(Exception):
    def __init__(self, code, msg):
        super().__init__('oracle error %d: %s' % (code, msg))
        self.code = code


def _check(rc):
    if rc != 0:
        raise OracleError(rc, lib().oracle_last_error().decode('latin-1'))


def _blob(names):
    off = np.zeros(len(names) + 1, dtype=np.int64)
    if len(names):
        off[1:] = np.cumsum([len(n) for n in names])
    return b''.join(names), off


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def py2_str_float(x):
    buf = C.create_string_buffer(64)
    lib().oracle_py2_str_float(C.c_double(x), buf, C.c_size_t(64))
    return buf.value.decode()


def round12(x):
    return lib().oracle_round12(float(x))


def make_links(bed, names, lengths, has_re, re_left, re_right, avoid=(), nthreads=1):
    """oracle_make_links.  `names` list of bytes; returns dict of numpy arrays
    (rows sorted by (first name, second name, ends)) with table indices."""
    L = lib()
    blob, off = _blob(names)
    lengths = np.ascontiguousarray(lengths, dtype=np.float64)
    has_re = np.ascontiguousarray(has_re, dtype=np.uint8)
    re_left = np.ascontiguousarray(re_left, dtype=np.int64)
    re_right = np.ascontiguousarray(re_right, dtype=np.int64)
    flat = [x for pr in avoid for x in pr]
    ablob, aoff = _blob(flat)
    out = _Links()
    bedbuf = bed if isinstance(bed, (bytes, bytearray)) else bytes(bed)
    rc = L.oracle_make_links(C.c_char_p(bedbuf), C.c_size_t(len(bedbuf)), C.c_int32(len(names)),
                             C.c_char_p(blob), _p(off, C.c_int64), _p(lengths, C.c_double), _p(has_re, C.c_uint8),
                             _p(re_left, C.c_int64), _p(re_right, C.c_int64),
                             C.c_int32(len(avoid)), C.c_char_p(ablob), _p(aoff, C.c_int64),
                             C.c_int(nthreads), C.byref(out))
    try:
        _check(rc)
        n = out.n
        res = dict(
            first=np.ctypeslib.as_array(out.first, (n,)).copy() if n else np.zeros(0, np.int32),
            second=np.ctypeslib.as_array(out.second, (n,)).copy() if n else np.zeros(0, np.int32),
            ends=np.ctypeslib.as_array(out.ends, (n,)).copy() if n else np.zeros(0, np.uint8),
            count=np.ctypeslib.as_array(out.count, (n,)).copy() if n else np.zeros(0, np.int64),
            norm=np.ctypeslib.as_array(out.norm, (n,)).copy() if n else np.zeros(0, np.float64),
            score=np.ctypeslib.as_array(out.score, (n,)).copy() if n else np.zeros(0, np.float64),
            score_is_int0=np.ctypeslib.as_array(out.score_is_int0, (n,)).copy() if n else np.zeros(0, np.uint8),
            nlines=out.nlines, npairs_counted=out.npairs_counted)
    finally:
        L.oracle_links_free(C.byref(out))
    return res


def link_lines(res, names):
    """contig_links_iteration_N text lines (make_links.py:169) for a make_links() result."""
    out = []
    for i in range(len(res['count'])):
        a, b, e = names[res['first'][i]], names[res['second'][i]], int(res['ends'][i])
        s = '0' if res['score_is_int0'][i] else py2_str_float(res['score'][i])
        out.append('%s:%s\t%s:%s\t%s\t%d\n' % (a.decode('latin-1'), 'E' if e & 2 else 'B',
                                             b.decode('latin-1'), 'E' if e & 1 else 'B', s, res['count'][i]))
    return out


def scaled_scores(u, v, w, nnodes):
    L = lib()
    u = np.ascontiguousarray(u, dtype=np.int32)
    v = np.ascontiguousarray(v, dtype=np.int32)
    w = np.ascontiguousarray(w, dtype=np.float64)
    n = len(u)
    best = np.zeros(n, np.float64)
    isint = np.zeros(n, np.uint8)
    ratio = np.zeros(n, np.float64)
    _check(L.oracle_scaled_scores(C.c_int64(n), C.c_int32(nnodes), _p(u, C.c_int32), _p(v, C.c_int32),
                                  _p(w, C.c_double), _p(best, C.c_double), _p(isint, C.c_uint8), _p(ratio, C.c_double)))
    return best, isint, ratio


def scaled_lines_from_link_lines(lines):
    """fast_scaled_scores.py as a text->text function: contig_links lines in,
    contig_links_scaled lines out (same order)."""
    nodes = {}
    u, v, w, links = [], [], [], []
    for ln in lines:
        a = ln.strip().split()
        for k in (0, 1):
            if a[k] not in nodes:
                nodes[a[k]] = len(nodes)
        u.append(nodes[a[0]])
        v.append(nodes[a[1]])
        w.append(float(a[2]))
        links.append(int(a[3]))
    best, isint, ratio = scaled_scores(u, v, w, max(len(nodes), 1))
    inv = list(nodes.keys())
    out = []
    for i in range(len(u)):
        b = '1' if isint[i] else py2_str_float(best[i])
        out.append('%s\t%s\t%s\t%s\t%s\t%d\tFalse\tFalse\n' % (inv[u[i]], inv[v[i]], py2_str_float(w[i]), b,
                                                                 py2_str_float(ratio[i]), links[i]))
    return out


def re_sites(fasta, enzyme):
    """Returns (names list[bytes], left, right, seqlen) in first-appearance order."""
    L = lib()
    out = _Re()
    buf = fasta if isinstance(fasta, (bytes, bytearray)) else bytes(fasta)
    rc = L.oracle_re_sites(C.c_char_p(buf), C.c_size_t(len(buf)), C.c_char_p(enzyme.encode()), C.byref(out))
    try:
        _check(rc)
        n = out.n
        off = np.ctypeslib.as_array(out.name_off, (n + 1,)).copy()
        blob = C.string_at(out.names_blob, int(off[-1])) if n else b''
        names = [blob[off[i]:off[i + 1]] for i in range(n)]
        left = np.ctypeslib.as_array(out.left, (n,)).copy() if n else np.zeros(0, np.int64)
        right = np.ctypeslib.as_array(out.right, (n,)).copy() if n else np.zeros(0, np.int64)
        seqlen = np.ctypeslib.as_array(out.seqlen, (n,)).copy() if n else np.zeros(0, np.int64)
    finally:
        L.oracle_re_free(C.byref(out))
    return names, left, right, seqlen


def cov_start(bed, names, lengths):
    """Returns list of (contig index, consensus position) -- break_contigs_start."""
    L = lib()
    blob, off = _blob(names)
    lengths = np.ascontiguousarray(lengths, dtype=np.int64)
    oc = np.zeros(len(names) + 1, np.int32)
    op = np.zeros(len(names) + 1, np.int64)
    on = C.c_int64(0)
    _check(L.oracle_cov_start(C.c_char_p(bed), C.c_size_t(len(bed)), C.c_int32(len(names)), C.c_char_p(blob),
                              _p(off, C.c_int64), _p(lengths, C.c_int64), _p(oc, C.c_int32), _p(op, C.c_int64),
                              C.byref(on)))
    return [(int(oc[i]), int(op[i])) for i in range(on.value)]


def cov_joins(bed, names, lengths, breakpoints):
    """breakpoints: dict name(bytes) -> list of positions (the breakpoints file).
    Returns (rows, total, suspicious); rows = list of (scaffold index, [suspicious positions])."""
    L = lib()
    blob, off = _blob(names)
    lengths = np.ascontiguousarray(lengths, dtype=np.int64)
    nc = len(names)
    inb = np.zeros(nc + 1, np.uint8)
    joff = np.zeros(nc + 1, np.int64)
    jpos = []
    for i, nm in enumerate(names):
        if nm in breakpoints:
            inb[i] = 1
            jpos.extend(breakpoints[nm])
        joff[i + 1] = len(jpos)
    jp = np.asarray(jpos + [0], dtype=np.int64)
    oscaf = np.zeros(nc + 1, np.int32)
    osoff = np.zeros(nc + 2, np.int64)
    ospos = np.zeros(len(jpos) + 1, np.int64)
    onsc, tot, sus = C.c_int64(0), C.c_int64(0), C.c_int64(0)
    _check(L.oracle_cov_joins(C.c_char_p(bed), C.c_size_t(len(bed)), C.c_int32(nc), C.c_char_p(blob),
                              _p(off, C.c_int64), _p(lengths, C.c_int64), _p(inb, C.c_uint8), _p(joff, C.c_int64),
                              _p(jp, C.c_int64), _p(oscaf, C.c_int32), _p(osoff, C.c_int64), _p(ospos, C.c_int64),
                              C.byref(onsc), C.byref(tot), C.byref(sus)))
    rows = []
    for k in range(onsc.value):
        rows.append((int(oscaf[k]), [int(x) for x in ospos[osoff[k]:osoff[k + 1]]]))
    return rows, tot.value, sus.value


def cov_analyse(diff, seed, lo, hi):
    """Prefix + total + the 11 Kadane windows of one segment (oracle_cov_analyse)."""
    L = lib()
    diff = np.ascontiguousarray(diff, dtype=np.int32)
    s = C.c_int64(0)
    st = np.zeros(11, np.int64)
    en = np.zeros(11, np.int64)
    _check(L.oracle_cov_analyse(_p(diff, C.c_int32), C.c_int64(len(diff)), C.c_int32(seed), C.c_int64(lo), C.c_int64(hi),
                                C.byref(s), _p(st, C.c_int64), _p(en, C.c_int64)))
    return s.value, st, en


# ---------------------------------------------------------------------------
# file-level helpers: the same argv/file contract as the reference programs
# ---------------------------------------------------------------------------
def read_lengths(path):
    """scaffold_length_iteration_N -> (names, float lengths); later duplicate wins."""
    d = {}
    with open(path, 'rb') as f:
        for ln in f:
            a = ln.split()
            d[a[0]] = float(a[1])
    return list(d.keys()), np.array(list(d.values()), dtype=np.float64)


def make_links_files(mapping, directory, iteration):
    """make_links.py -b mapping -d directory -i iteration, output sorted (see C header)."""
    it = str(iteration)
    names, lengths = read_lengths(directory + '/scaffold_length_iteration_' + it)
    idx = {n: i for i, n in enumerate(names)}
    has_re = np.zeros(len(names), np.uint8)
    rl = np.zeros(len(names), np.int64)
    rr = np.zeros(len(names), np.int64)
    with open(directory + '/re_counts_iteration_' + it, 'rb') as f:
        for ln in f:
            a = ln.split()
            if a[0] in idx:
                i = idx[a[0]]
                has_re[i], rl[i], rr[i] = 1, int(a[1]), int(a[2])
    avoid = []
    if int(it) > 1:
        with open(directory + '/links_avoid_iteration_' + it, 'rb') as f:
            for ln in f:
                a = ln.split()
                avoid.append((a[0], a[1]))
    with open(mapping, 'rb') as f:
        bed = f.read()
    res = make_links(bed, names, lengths, has_re, rl, rr, avoid)
    lines = link_lines(res, names)
    with open(directory + '/contig_links_iteration_' + it, 'w', encoding='latin-1') as f:
        f.writelines(lines)
    return lines


def fast_scaled_scores_files(directory, iteration):
    it = str(iteration)
    with open(directory + '/contig_links_iteration_' + it, 'r', encoding='latin-1') as f:
        lines = f.readlines()
    out = scaled_lines_from_link_lines(lines)
    with open(directory + '/contig_links_scaled_iteration_' + it, 'w', encoding='latin-1') as f:
        f.writelines(out)
    return out


def ref_binary(name):
    """Path of a compiled UNMODIFIED reference program in oracle/_ref, or None."""
    p = os.path.join(REF_DIR, name)
    return p if os.path.exists(p) else None
