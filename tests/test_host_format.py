"""The two host-side file formatters of the library (salsa_format_links, salsa_format_scaled; no GPU work): a large call,
which is cut into one piece per host thread, must give byte for byte what the same edges give in small calls (one
thread each), and must name the first edge it cannot write."""
import ctypes as C

import numpy as np
import pytest


@pytest.fixture(scope='module')
def L():
    from salsa_b200 import _lib
    return _lib.lib()


def _names(nn):
    names = [b'ctg%06d:%s' % (i // 2, b'BE'[i % 2:i % 2 + 1]) for i in range(nn)]
    off = np.zeros(nn + 1, np.int64)
    off[1:] = np.cumsum([len(x) for x in names])
    return b''.join(names), off


def _digits(rng, n):
    d = rng.integers(10 ** 11, 10 ** 12, n).astype(np.int64)
    d[rng.random(n) < 0.05] = 0
    return d, rng.integers(-8, 14, n).astype(np.int32)


def test_threaded_formatters_equal_piecewise_calls(L):
    rng = np.random.default_rng(3)
    nn, n = 3000, 150_000
    blob, off = _names(nn)
    u, v = rng.integers(0, nn, n).astype(np.int32), rng.integers(0, nn, n).astype(np.int32)
    (wd, wk), (bd, bk), (rd, rk) = _digits(rng, n), _digits(rng, n), _digits(rng, n)
    isint = (rng.random(n) < 0.1).astype(np.uint8)
    links = rng.integers(1, 10 ** 7, n).astype(np.int64)

    def scaled(lo, hi, cols=(u, v, wd, wk, bd, bk, isint, rd, rk, links)):
        m, cap = hi - lo, (hi - lo) * 200 + 1024
        out = np.empty(cap, np.uint8)
        a = [np.ascontiguousarray(x[lo:hi]) for x in cols]
        nb = L.salsa_format_scaled(blob, off.ctypes.data, *[x.ctypes.data for x in a], m, out.ctypes.data, cap)
        return nb, out[:max(nb, 0)].tobytes()

    nb, whole = scaled(0, n)
    assert nb > 0
    assert whole == b''.join(scaled(lo, min(n, lo + 20000))[1] for lo in range(0, n, 20000))
    assert whole.count(b'\n') == n and whole.endswith(b'\tFalse\tFalse\n')

    first, second = rng.integers(0, nn, n).astype(np.int64), rng.integers(0, nn, n).astype(np.int64)
    ends = rng.integers(0, 4, n).astype(np.uint8)
    cnt = rng.integers(1, 10 ** 6, n).astype(np.int64)
    isz = (rng.random(n) < 0.02).astype(np.uint8)

    def linksf(lo, hi):
        m, cap = hi - lo, (hi - lo) * 200 + 1024
        out = np.empty(cap, np.uint8)
        a = [np.ascontiguousarray(x[lo:hi]) for x in (first, second, ends, cnt, wd, wk, isz)]
        nb = L.salsa_format_links(blob, off.ctypes.data, *[x.ctypes.data for x in a], m, out.ctypes.data, cap)
        assert nb >= 0
        return out[:nb].tobytes()

    assert linksf(0, n) == b''.join(linksf(lo, min(n, lo + 20000)) for lo in range(0, n, 20000))

    # a value no digits exist for (salsa_b200.h: INT64_MIN): refused, naming the edge -- the lowest one
    bad = wd.copy()
    bad[[140_000, 70_001]] = np.iinfo(np.int64).min
    nb, _ = scaled(0, n, cols=(u, v, bad, wk, bd, bk, isint, rd, rk, links))
    assert nb == -1 and b'edge 70001 ' in L.salsa_last_error()
    # a buffer that is too small is refused as well
    out = np.empty(1000, np.uint8)
    a = [np.ascontiguousarray(x) for x in (u, v, wd, wk, bd, bk, isint, rd, rk, links)]
    assert L.salsa_format_scaled(blob, off.ctypes.data, *[x.ctypes.data for x in a], n, out.ctypes.data, 1000) == -1
    assert b'too small' in L.salsa_last_error()
