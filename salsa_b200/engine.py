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

    def scan_mode(self, mode):
        """Special-byte test of the link scan: 0 = a sample of the text decides (default), 1 = exact, 2 = cheap
        superset.  Results never depend on it (include/salsa_b200.h: salsa_ctx_scan_mode)."""
        check(self.L.salsa_ctx_scan_mode(self.ctx, int(mode)))

    def timers(self, enable=None, reset=False):
        """Device time of the BED link scans launched through this engine (salsa_timers): (ms, launches).
        timers(enable=True) switches the event bracketing on first."""
        if enable is not None:
            check(self.L.salsa_timers_enable(self.ctx, 1 if enable else 0))
        ms = (C.c_double * 4)()
        nl = (C.c_uint64 * 4)()
        check(self.L.salsa_timers(self.ctx, ms, nl, 4, 1 if reset else 0))
        return float(ms[0]), int(nl[0])

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
        self.involved = set(x for a, b in avoid if a != b for x in (a, b))
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
            cap = 1 << 10
            while cap < (sparse_capacity or (1 << 24)):
                cap <<= 1
            keys = torch.empty(cap, dtype=torch.int64, device=self.device)
            cnts = torch.empty(cap, dtype=torch.int64, device=self.device)
            check(self.L.salsa_sparse_reset(keys.data_ptr(), cnts.data_ptr(), cap, st))
            self.sparse = (keys, cnts, cap)
            self.table = None

    def ingest_links(self, ingest, mode='auto', sparse_capacity=None):
        """begin_links(mode) + ingest() + stats_read().  A sparse table that turns out too small is not an error of
        the input -- the reference's dict simply grows (make_links.py:140-144) -- so on SALSA_DATA_TABLE_FULL the table
        is made four times larger and the BED is ingested again, until it fits or the GPU has no room for the table.
        `ingest` is a callable that feeds the whole BED (ingest_device / ingest_file calls)."""
        cap = sparse_capacity
        while True:
            self.begin_links(mode, sparse_capacity=cap)
            ingest()
            try:
                return self.stats_read()
            except DataError as e:
                if e.code != _lib.DATA_TABLE_FULL or self.mode != 'sparse':
                    raise
                cap = self.sparse[2] * 4
                free, _ = torch.cuda.mem_get_info(self.device)
                if cap * 16 > free + self.sparse[2] * 16:
                    raise MemoryError('sparse link table of %d slots does not fit the GPU' % cap)
                self.sparse = None

    def ingest_device(self, bed, base_offset=0):
        """bed: uint8 CUDA tensor whose first byte starts a line.  Asynchronous."""
        assert bed.dtype == torch.uint8 and bed.is_cuda and bed.is_contiguous()
        st = torch.cuda.current_stream().cuda_stream
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
        if self.mode != 'dense':
            raise NotImplementedError('host streaming is wired for the dense table only')
        if stage is None:
            stage = torch.empty(2 * chunk_bytes, dtype=torch.uint8, device=self.device)
        chunk_bytes = stage.numel() // 2 // 16 * 16
        check(self.L.salsa_bed_ingest_dense_host(self.ctx, ptr, n, stage.data_ptr(), chunk_bytes, self.table.data_ptr(),
                                                 self.stats.data_ptr(), torch.cuda.current_stream().cuda_stream))

    def ingest_file(self, path, chunk_bytes=256 << 20, whole=None):
        """Ingest a BED file from disk (what make_links.py does with -b).  The file
        is memory-mapped and streamed in line-aligned chunks that overlap by one
        line (see salsa_bed_ingest_dense in include/salsa_b200.h)."""
        import os
        size = os.path.getsize(path)
        if size == 0:
            return
        # A BED that fits the GPU goes there in one piece, through four readers and pinned buffers (a pageable copy out of
        # the mapping runs at a fraction of the host link); only a larger one is cut into chunks.  With an avoid list this
        # is also what makes every file ingestible: a chunk may only start on a line that can never be skipped
        # (make_links.py:72-74), and a BED whose every scaffold is in some avoid pair has no such line, while the kernel
        # itself resolves chains of any length.
        free, _ = torch.cuda.mem_get_info(self.device)
        if whole is None:
            whole = size + (256 << 20) < free             # (whole=False: the tests' way to the chunked path)
        if whole:
            from .reproject import _stream_in
            dev = _stream_in(os.path.realpath(path), size, self.device)
            self.ingest_device(dev)
            torch.cuda.current_stream().synchronize()
            return
        mm = np.memmap(path, dtype=np.uint8, mode='r')
        if self.mode == 'dense':
            self.ingest_host(mm, chunk_bytes=min(chunk_bytes, 64 << 20))
            return
        pos = 0
        while pos < size:
            end = min(pos + chunk_bytes, size)
            nxt = size
            if end < size:
                blk = mm[pos:end]
                nl = np.flatnonzero(blk[::-1][:4 * 1024 * 1024] == 10)      # newlines near the chunk's end
                if len(nl) < 2:
                    raise ValueError('BED line longer than the staging chunk near byte %d' % pos)
                k = 0
                if self.navoid:
                    # the next chunk must start on a line that can never be skipped: contig in no avoid pair
                    while True:
                        if k + 1 >= len(nl):
                            raise ValueError('no line outside the avoid list near byte %d to split the BED at' % pos)
                        ls = len(blk) - 1 - int(nl[k + 1]) + 1                # start of the last complete line
                        tok = bytes(blk[ls:ls + 1024]).split(None, 1)
                        if not tok or tok[0] not in self.involved:
                            break
                        k += 1
                end = pos + (len(blk) - 1 - int(nl[k])) + 1
                nxt = pos + (len(blk) - 1 - int(nl[k + 1])) + 1
            dev = torch.from_numpy(np.array(mm[pos:end])).to(self.device)      # a writable copy (torch refuses read-only maps)
            self.ingest_device(dev, base_offset=pos)
            torch.cuda.current_stream().synchronize()
            pos = nxt

    def stats_read(self, raise_on_error=True):
        s = IngestStats()
        check(self.L.salsa_ingest_stats_read(self.stats.data_ptr(), C.byref(s), torch.cuda.current_stream().cuda_stream))
        if raise_on_error and s.err_code:
            raise DataError(s.err_code, s.err_offset)
        return s

    def _compact(self):
        """(keys, counts, m): compacted, key-sorted link list on the device."""
        st = torch.cuda.current_stream().cuda_stream
        if self.mode == 'dense':
            kptr, cnt, ncell = None, self.table, self.table.numel()
        else:
            keys, cnt, ncell = self.sparse
            kptr = keys.data_ptr()
        scratch = torch.empty(self.L.salsa_links_compact_scratch_bytes(ncell) // 8 + 1, dtype=torch.int64, device=self.device)
        n = C.c_uint64(0)
        # one pass: there are at most as many links as counted pairs (and as cells)
        bound = min(ncell, self.stats_read(raise_on_error=False).pairs)
        ok = torch.empty(max(bound, 1), dtype=torch.int64, device=self.device)
        oc = torch.empty(max(bound, 1), dtype=torch.int64, device=self.device)
        check(self.L.salsa_links_compact(kptr, cnt.data_ptr(), ncell, ok.data_ptr(), oc.data_ptr(), bound, scratch.data_ptr(),
                                         C.byref(n), st))
        m = n.value
        if self.mode != 'dense' and m:
            ok2, oc2 = torch.empty_like(ok), torch.empty_like(oc)
            tb = C.c_size_t(0)
            check(self.L.salsa_sort_pairs_u64(ok.data_ptr(), oc.data_ptr(), ok2.data_ptr(), oc2.data_ptr(), m, None, C.byref(tb), st))
            tmp = torch.empty(tb.value + 16, dtype=torch.uint8, device=self.device)
            check(self.L.salsa_sort_pairs_u64(ok.data_ptr(), oc.data_ptr(), ok2.data_ptr(), oc2.data_ptr(), m, tmp.data_ptr(),
                                              C.byref(tb), st))
            ok, oc = ok2, oc2
        return ok, oc, m

    def _decode(self, ok, m):
        keys = ok[:m]                                  # < 2^63: n*n*4 cells
        pair = keys >> 2
        first = torch.div(pair, max(self.n, 1), rounding_mode='floor')
        second = pair - first * max(self.n, 1)
        return first.cpu().numpy(), second.cpu().numpy(), (keys & 3).to(torch.uint8).cpu().numpy()

    def links(self):
        """Compacted, key-sorted link list as numpy arrays (first, second, ends, count)."""
        ok, oc, m = self._compact()
        first, second, ends = self._decode(ok, m)
        return first, second, ends, oc[:m].cpu().numpy()

    def score_links(self):
        """make_links.py:161-169 on the device: per link the score, its str(float)
        digits and the value float(str(score)) the next program sees."""
        ok, oc, m = self._compact()
        dev, st = self.device, torch.cuda.current_stream().cuda_stream
        mm = max(m, 1)
        rl = torch.from_numpy(self.re_left).to(dev)
        rr = torch.from_numpy(self.re_right).to(dev)
        score = torch.empty(mm, dtype=torch.float64, device=dev)
        w12 = torch.empty(mm, dtype=torch.float64, device=dev)
        int0 = torch.empty(mm, dtype=torch.uint8, device=dev)
        digits = torch.empty(mm, dtype=torch.int64, device=dev)
        exp10 = torch.empty(mm, dtype=torch.int32, device=dev)
        u = torch.empty(mm, dtype=torch.int32, device=dev)
        v = torch.empty(mm, dtype=torch.int32, device=dev)
        check(self.L.salsa_links_score(ok.data_ptr(), oc.data_ptr(), m, self.n, rl.data_ptr(), rr.data_ptr(), score.data_ptr(),
                                       int0.data_ptr(), digits.data_ptr(), exp10.data_ptr(), w12.data_ptr(), u.data_ptr(),
                                       v.data_ptr(), st))
        first, second, ends = self._decode(ok, m)
        return dict(first=first, second=second, ends=ends, count=oc[:m].cpu().numpy(), score=score[:m].cpu().numpy(),
                    is_int0=int0[:m].cpu().numpy(), digits=digits[:m].cpu().numpy(), exp10=exp10[:m].cpu().numpy(),
                    w12=w12[:m], u=u[:m], v=v[:m], n=m)

    def link_text(self, sc):
        """contig_links_iteration_N content for a score_links() result (bytes)."""
        blob, off = _blob(self.names)
        cap = int(sc['n']) * 96 + 2 * int(off[-1] if len(off) else 0) + sum(0 for _ in ()) + 1024
        cap += int((np.diff(off)[sc['first']].sum() + np.diff(off)[sc['second']].sum())) if sc['n'] else 0
        out = np.empty(cap, dtype=np.uint8)              # (create_string_buffer would zero it, .raw would copy all of it)
        f = np.ascontiguousarray(sc['first'], np.int64); s = np.ascontiguousarray(sc['second'], np.int64)
        e = np.ascontiguousarray(sc['ends'], np.uint8); c = np.ascontiguousarray(sc['count'], np.int64)
        d = np.ascontiguousarray(sc['digits'], np.int64); k = np.ascontiguousarray(sc['exp10'], np.int32)
        z = np.ascontiguousarray(sc['is_int0'], np.uint8)
        nb = self.L.salsa_format_links(blob, off.ctypes.data, f.ctypes.data, s.ctypes.data, e.ctypes.data, c.ctypes.data,
                                       d.ctypes.data, k.ctypes.data, z.ctypes.data, int(sc['n']), out.ctypes.data, cap)
        if nb < 0:
            check(-2)
        return out[:nb].tobytes()

    def g12(self, x):
        """(digits, exp10, w12) of Python-2 str(float) for a float64 array (device kernel)."""
        x = torch.as_tensor(np.ascontiguousarray(x, np.float64)).to(self.device)
        n = x.numel()
        d = torch.empty(max(n, 1), dtype=torch.int64, device=self.device)
        k = torch.empty(max(n, 1), dtype=torch.int32, device=self.device)
        w = torch.empty(max(n, 1), dtype=torch.float64, device=self.device)
        check(self.L.salsa_g12(x.data_ptr(), n, d.data_ptr(), k.data_ptr(), w.data_ptr(), torch.cuda.current_stream().cuda_stream))
        return d[:n].cpu().numpy(), k[:n].cpu().numpy(), w[:n].cpu().numpy()

    def scaled_scores(self, u, v, w, nnodes):
        """fast_scaled_scores.py:35-89 on the device.  u, v int32 node ids, w float64
        (device tensors or numpy).  Returns dict of numpy arrays."""
        dev, st = self.device, torch.cuda.current_stream().cuda_stream
        u = torch.as_tensor(u).to(dev, torch.int32).contiguous()
        v = torch.as_tensor(v).to(dev, torch.int32).contiguous()
        w = torch.as_tensor(w).to(dev, torch.float64).contiguous()
        m = u.numel()
        mm = max(m, 1)
        scratch = torch.empty(self.L.salsa_scaled_scratch_bytes(nnodes) // 8 + 1, dtype=torch.int64, device=dev)
        best = torch.empty(mm, dtype=torch.float64, device=dev)
        int1 = torch.empty(mm, dtype=torch.uint8, device=dev)
        ratio = torch.empty(mm, dtype=torch.float64, device=dev)
        rd = torch.empty(mm, dtype=torch.int64, device=dev); rk = torch.empty(mm, dtype=torch.int32, device=dev)
        bd = torch.empty(mm, dtype=torch.int64, device=dev); bk = torch.empty(mm, dtype=torch.int32, device=dev)
        check(self.L.salsa_scaled_scores(u.data_ptr(), v.data_ptr(), w.data_ptr(), m, nnodes, scratch.data_ptr(), best.data_ptr(),
                                         int1.data_ptr(), ratio.data_ptr(), rd.data_ptr(), rk.data_ptr(), bd.data_ptr(),
                                         bk.data_ptr(), st))
        g = lambda t: t[:m].cpu().numpy()   # noqa: E731
        return dict(best_alt=g(best), is_int1=g(int1), ratio=g(ratio), r_digits=g(rd), r_exp10=g(rk), b_digits=g(bd), b_exp10=g(bk))
