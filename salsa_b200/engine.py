"""Host-side mirror of the reference's hot-path programs on top of the C ABI.

PyTorch is used only to own device buffers and streams; every computation on
the path is a kernel of libsalsa_b200.so.  Nothing here falls back to the CPU:
without the shared library or without a CUDA device the constructors raise.
"""
import ctypes as C
import math

import numpy as np
import torch

from . import _lib
from ._lib import DataError, IngestStats, SalsaError, check  # noqa: F401

HAS_LEN, HAS_RE = 1, 2
DENSE_MAX_BYTES = 8 << 30      # above this the link table goes sparse


def _blob(names):
    off = np.zeros(len(names) + 1, dtype=np.int64)
    if len(names):
        off[1:] = np.cumsum([len(n) for n in names])
    return b''.join(names), off


def length_floor(x):
    """floor(float(text)) with the checks that keep the integer B/E rule exact
    (make_links.py:37-41,104-139; DESIGN.md 'end classification')."""
    v = float(x)
    if not math.isfinite(v) or abs(v) >= 2 ** 53:
        raise ValueError('unsupported contig length %r' % (x,))
    if 0 < v < 1:
        raise ValueError('unsupported fractional contig length %r' % (x,))
    return int(math.floor(v))


class Engine:
    """One per (process, GPU).  Owns the salsa_ctx and the link table."""

    def __init__(self, device=0):
        self.L = _lib.lib()
        if not torch.cuda.is_available():
            raise RuntimeError('salsa_b200 needs a CUDA device (B200); there is no CPU fallback')
        self.device = torch.device('cuda', device)
        torch.cuda.set_device(self.device)
        h = C.c_void_p()
        check(self.L.salsa_ctx_create(device, C.byref(h)))
        self.ctx = h
        self.names = []
        self.n = 0
        self.table = None
        self.sparse = None
        self.stats = torch.zeros(4, dtype=torch.int64, device=self.device)

    def close(self):
        if self.ctx:
            self.L.salsa_ctx_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ contigs
    def set_contigs(self, names, len_floor, re_left=None, re_right=None, has_re=None, avoid=()):
        """names: list[bytes] in any order (duplicates: the later entry wins, like
        the reference's dict assignment).  avoid: iterable of (bytes, bytes)."""
        d = {}
        for i, nm in enumerate(names):
            d[nm] = i
        extra = []
        for a, b in avoid:
            for x in (a, b):
                if x not in d and x not in extra:
                    extra.append(x)
        allnames = sorted(list(d.keys()) + extra)          # bytes order == Python-2 str order
        src = np.array([d.get(nm, -1) for nm in allnames], dtype=np.int64)
        n = len(allnames)
        lf = np.zeros(n, dtype=np.int64)
        fl = np.zeros(n, dtype=np.uint32)
        rl = np.zeros(n, dtype=np.int64)
        rr = np.zeros(n, dtype=np.int64)
        known = src >= 0
        len_floor = np.asarray(len_floor, dtype=np.int64)
        lf[known] = len_floor[src[known]]
        fl[known] |= HAS_LEN
        if re_left is not None:
            hr = np.ones(len(names), dtype=bool) if has_re is None else np.asarray(has_re).astype(bool)
            rl[known] = np.asarray(re_left, dtype=np.int64)[src[known]]
            rr[known] = np.asarray(re_right, dtype=np.int64)[src[known]]
            fl[known] |= np.where(hr[src[known]], HAS_RE, 0).astype(np.uint32)
        blob, off = _blob(allnames)
        check(self.L.salsa_names_build(self.ctx, blob, off.ctypes.data, n, lf.ctypes.data, fl.ctypes.data))
        self.names, self.n = allnames, n
        self.len_floor, self.flags, self.re_left, self.re_right = lf, fl, rl, rr
        idx = {nm: i for i, nm in enumerate(allnames)}
        pairs = np.array([[idx[a], idx[b]] for a, b in avoid], dtype=np.int32).reshape(-1, 2)
        check(self.L.salsa_avoid_set(self.ctx, pairs.ctypes.data if len(pairs) else None, len(pairs)))
        self.navoid = len(pairs)
        self.table = None
        self.sparse = None
        return idx

    # ------------------------------------------------------------------ table
    def begin_links(self, mode='auto', sparse_capacity=None):
        """Zero the link table.  dense: uint64 [n*n*4]; sparse: open-addressing table."""
        ncell = self.n * self.n * 4
        if mode == 'auto':
            mode = 'dense' if ncell * 8 <= DENSE_MAX_BYTES else 'sparse'
        self.mode = mode
        st = torch.cuda.current_stream().cuda_stream
        check(self.L.salsa_ingest_stats_reset(self.stats.data_ptr(), st))
        if mode == 'dense':
            self.table = torch.zeros(ncell, dtype=torch.int64, device=self.device)
            self.sparse = None
        else:
            cap = 1 << 20
            while cap < (sparse_capacity or (1 << 24)):
                cap <<= 1
            keys = torch.empty(cap, dtype=torch.int64, device=self.device)
            cnts = torch.empty(cap, dtype=torch.int64, device=self.device)
            check(self.L.salsa_sparse_reset(keys.data_ptr(), cnts.data_ptr(), cap, st))
            self.sparse = (keys, cnts, cap)
            self.table = None

    def ingest_device(self, bed, base_offset=0):
        """bed: uint8 CUDA tensor whose first byte starts a line.  Asynchronous."""
        assert bed.dtype == torch.uint8 and bed.is_cuda and bed.is_contiguous()
        st = torch.cuda.current_stream().cuda_stream
        if self.navoid:
            raise NotImplementedError('avoid-list ingest')
        if self.mode == 'dense':
            check(self.L.salsa_bed_ingest_dense(self.ctx, bed.data_ptr(), bed.numel(), base_offset,
                                                self.table.data_ptr(), self.stats.data_ptr(), st))
        else:
            keys, cnts, cap = self.sparse
            check(self.L.salsa_bed_ingest_sparse(self.ctx, bed.data_ptr(), bed.numel(), base_offset, keys.data_ptr(),
                                                 cnts.data_ptr(), cap, self.stats.data_ptr(), st))

    def ingest_host(self, buf, nbytes=None, stage=None, chunk_bytes=64 << 20):
        """buf: host memory (numpy uint8 array / pinned torch tensor).  Streams it
        through a double-buffered device staging area.  Synchronous."""
        if isinstance(buf, torch.Tensor):
            ptr, n = buf.data_ptr(), buf.numel()
        else:
            ptr, n = buf.ctypes.data, buf.size
        if nbytes is not None:
            n = nbytes
        if self.mode != 'dense' or self.navoid:
            raise NotImplementedError('host streaming is wired for the dense table only')
        if stage is None:
            stage = torch.empty(2 * chunk_bytes, dtype=torch.uint8, device=self.device)
        chunk_bytes = stage.numel() // 2 // 16 * 16
        torch.cuda.current_stream().synchronize()
        check(self.L.salsa_bed_ingest_dense_host(self.ctx, ptr, n, stage.data_ptr(), chunk_bytes,
                                                 self.table.data_ptr(), self.stats.data_ptr()))

    def stats_read(self, raise_on_error=True):
        s = IngestStats()
        check(self.L.salsa_ingest_stats_read(self.stats.data_ptr(), C.byref(s), torch.cuda.current_stream().cuda_stream))
        if raise_on_error and s.err_code:
            raise DataError(s.err_code, s.err_offset)
        return s

    def links(self):
        """Compacted, key-sorted link list as numpy arrays (first, second, ends, count)."""
        st = torch.cuda.current_stream().cuda_stream
        if self.mode == 'dense':
            kptr, cnt, ncell = None, self.table, self.table.numel()
        else:
            keys, cnt, ncell = self.sparse
            kptr = keys.data_ptr()
        scratch = torch.empty(self.L.salsa_links_compact_scratch_bytes(ncell) // 8 + 1, dtype=torch.int64, device=self.device)
        n = C.c_uint64(0)
        check(self.L.salsa_links_compact(kptr, cnt.data_ptr(), ncell, None, None, 0, scratch.data_ptr(), C.byref(n), st))
        m = n.value
        ok = torch.empty(max(m, 1), dtype=torch.int64, device=self.device)
        oc = torch.empty(max(m, 1), dtype=torch.int64, device=self.device)
        check(self.L.salsa_links_compact(kptr, cnt.data_ptr(), ncell, ok.data_ptr(), oc.data_ptr(), m, scratch.data_ptr(),
                                         C.byref(n), st))
        if self.mode != 'dense' and m:
            ok2, oc2 = torch.empty_like(ok), torch.empty_like(oc)
            tb = C.c_size_t(0)
            check(self.L.salsa_sort_pairs_u64(ok.data_ptr(), oc.data_ptr(), ok2.data_ptr(), oc2.data_ptr(), m, None, C.byref(tb), st))
            tmp = torch.empty(tb.value + 16, dtype=torch.uint8, device=self.device)
            check(self.L.salsa_sort_pairs_u64(ok.data_ptr(), oc.data_ptr(), ok2.data_ptr(), oc2.data_ptr(), m, tmp.data_ptr(),
                                              C.byref(tb), st))
            ok, oc = ok2, oc2
        keys = ok[:m].cpu().numpy().astype(np.uint64)
        counts = oc[:m].cpu().numpy()
        ends = (keys & np.uint64(3)).astype(np.uint8)
        pair = keys >> np.uint64(2)
        first = (pair // np.uint64(max(self.n, 1))).astype(np.int64)
        second = (pair % np.uint64(max(self.n, 1))).astype(np.int64)
        return first, second, ends, counts
