"""GPU BED pass of update_bed: alignment_iteration_1.bed -> alignment_iteration_N.bed in
scaffold coordinates (layout_unitigs.py:816-886, refactor_breaks.py:133-203), on top of
salsa_reproject_candidates / salsa_sort_u64 / salsa_reproject_emit.

The iteration-1 BED is read from disk once per process and then stays in HBM (every
iteration re-projects it twice); the BED each pass produces also stays resident, so the
make_links.py / break_contigs steps that follow take it from there, and its file -- the
reference's resume contract -- is written behind their back (write_behind / flush).
"""
import ctypes as C
import os
import threading

import numpy as np
import torch

from ._lib import check
from .engine import _blob

CHUNK = 64 << 20
_RESIDENT = {}          # realpath -> (size, mtime_ns, uint8 CUDA tensor); size None while the file is still being written
_WRITERS = {}           # realpath -> _Writer still running (a writer takes itself out when its file is complete)
_FAILED = {}            # realpath -> exception of a writer that failed; raised by flush()
_LOCK = threading.Lock()


def resident_bed(path, device, load=True):
    """uint8 CUDA tensor holding the file: from the cache when the file's size and mtime still match (or when
    this process is still writing that very file), else streamed in from disk."""
    rp = os.path.realpath(path)
    hit = _RESIDENT.get(rp)
    if hit and hit[0] is None and rp in _WRITERS and hit[2].device == device:
        return hit[2]
    st = os.stat(rp)
    if hit and hit[0] == st.st_size and hit[1] == st.st_mtime_ns and hit[2].device == device:
        return hit[2]
    if not load:
        return None
    free, _ = torch.cuda.mem_get_info(device)
    if st.st_size * 3 > free:
        raise MemoryError('%s (%d bytes) is not kept on the GPU: it would not fit beside its re-projection' % (path, st.st_size))
    t = _stream_in(rp, st.st_size, device)
    _remember(rp, st.st_size, st.st_mtime_ns, t)
    return t


READERS = 4             # threads that read a large file into pinned memory side by side
_PINNED = {}            # (nbytes, slot) -> pinned staging buffer, kept for the life of the process (cudaHostAlloc is slow)


def _pinned(nbytes, slot):
    b = _PINNED.get((nbytes, slot))
    if b is None:
        b = _PINNED[(nbytes, slot)] = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    return b


def _stream_in(rp, size, device, src=None):
    """File (or, with src, a host uint8 numpy array such as a read-only np.memmap) -> device.  One reader's copy out of the page cache runs at ~7 GB/s, a tenth of what the host link takes, so
    files beyond a few chunks are read by READERS threads side by side (chunks dealt round robin; pread and the copies
    release the GIL), each through its own two pinned 32 MiB buffers and its own stream: the read of a thread's next
    chunk overlaps the copy of its last one."""
    t = torch.empty(size, dtype=torch.uint8, device=device)
    if size == 0:
        return t
    chunk = CHUNK // 2
    nchunk = (size + chunk - 1) // chunk
    nthr = max(1, min(READERS, nchunk // 2))
    errors = []
    start = torch.cuda.Event()
    start.record()                                            # t's allocation is ordered on the current stream

    def work(w):
        try:
            torch.cuda.set_device(device)
            stream = torch.cuda.Stream(device=device)
            stream.wait_event(start)
            pin = [_pinned(chunk, 2 * w), _pinned(chunk, 2 * w + 1)]
            ev = [torch.cuda.Event(), torch.cuda.Event()]
            fd = os.open(rp, os.O_RDONLY) if src is None else -1
            try:
                with torch.cuda.stream(stream):
                    for i, c in enumerate(range(w, nchunk, nthr)):
                        pos = c * chunk
                        n = min(chunk, size - pos)
                        if i >= 2:
                            ev[i & 1].synchronize()           # the copy that last used this buffer is done
                        view = memoryview(pin[i & 1].numpy())[:n]
                        got = 0
                        if src is not None:
                            np.copyto(pin[i & 1].numpy()[:n], src[pos:pos + n])      # (releases the GIL)
                            got = n
                        while got < n:
                            r = os.preadv(fd, [view[got:]], pos + got)
                            if not r:
                                raise IOError('%s changed size while being read' % rp)
                            got += r
                        t[pos:pos + n].copy_(pin[i & 1][:n], non_blocking=True)
                        ev[i & 1].record(stream)
                stream.synchronize()
            finally:
                if fd >= 0:
                    os.close(fd)
        except Exception as e:                                # raised in the caller's thread below
            errors.append(e)

    if nthr == 1:
        work(0)
    else:
        th = [threading.Thread(target=work, args=(w,)) for w in range(nthr)]
        for x in th:
            x.start()
        for x in th:
            x.join()
    if errors:
        raise errors[0]
    return t


class _Writer(threading.Thread):
    """Device tensor -> file in the background (pinned double buffer on a side stream), written to
    `path.part` and renamed when complete, so a reader -- or a resumed run -- never sees half a file."""

    def __init__(self, rp, t, ready):
        super().__init__(daemon=True)
        self.rp, self.t, self.ready, self.cancelled, self.error = rp, t, ready, False, None
        self.markers, self.done = [], False          # files to create once the BED is complete on disk (after_write)

    def run(self):
        try:
            t, size = self.t, self.t.numel()
            stream = torch.cuda.Stream(device=t.device)
            pin = [torch.empty(min(CHUNK, max(size, 1)), dtype=torch.uint8).pin_memory() for _ in range(2)]
            ev = [torch.cuda.Event(), torch.cuda.Event()]
            with torch.cuda.stream(stream):
                stream.wait_event(self.ready)
                spans = [(o, min(CHUNK, size - o)) for o in range(0, size, CHUNK)]
                with open(self.rp + '.part', 'wb', buffering=0) as f:
                    for k, (o, n) in enumerate(spans[:1]):
                        pin[0][:n].copy_(t[o:o + n], non_blocking=True)
                        ev[0].record(stream)
                    for k, (o, n) in enumerate(spans):
                        if self.cancelled:
                            break
                        if k + 1 < len(spans):
                            o2, n2 = spans[k + 1]
                            pin[(k + 1) & 1][:n2].copy_(t[o2:o2 + n2], non_blocking=True)
                            ev[(k + 1) & 1].record(stream)
                        ev[k & 1].synchronize()
                        f.write(memoryview(pin[k & 1].numpy()[:n]))
                stream.synchronize()
            if self.cancelled:
                os.remove(self.rp + '.part')
            else:
                os.replace(self.rp + '.part', self.rp)
        except Exception as e:           # reported by flush()
            self.error = e
        # The file is complete (or abandoned): the writer lets go of the tensor and of its place in _WRITERS, so that
        # _remember can drop the product once a newer one exists; markers appear only behind a complete file.
        with _LOCK:
            self.done = True
            if _WRITERS.get(self.rp) is self:
                del _WRITERS[self.rp]
            if self.error is not None:
                _FAILED[self.rp] = self.error
            elif not self.cancelled:
                hit = _RESIDENT.get(self.rp)
                if hit is not None and hit[0] is None and hit[2] is self.t:
                    st = os.stat(self.rp)
                    _RESIDENT[self.rp] = (st.st_size, st.st_mtime_ns, hit[2])
                for m in self.markers:
                    open(m, 'w').close()
            self.t = None


def write_behind(path, t):
    """Start writing t to `path` in the background; t stays the resident copy of that file.  A write of the
    same path that is still running is abandoned (its content is about to be replaced)."""
    rp = os.path.realpath(path)
    old = _WRITERS.get(rp)
    if old is not None:
        old.cancelled = True
        old.join()
    _FAILED.pop(rp, None)
    if os.path.lexists(rp):
        os.remove(rp)
    ready = torch.cuda.Event()
    ready.record()
    w = _Writer(rp, t, ready)
    with _LOCK:
        _WRITERS[rp] = w
        _remember(rp, None, None, t)
    w.start()


def after_write(path, marker):
    """Create the (empty) file `marker` once `path` is complete on disk: right away when nothing is writing it any more,
    else from the writer thread behind its rename -- a DONE marker never precedes the BED it vouches for."""
    rp = os.path.realpath(path)
    with _LOCK:
        w = _WRITERS.get(rp)
        if w is not None and not w.done:
            w.markers.append(marker)
            return
    err = _FAILED.pop(rp, None)
    if err is not None:
        raise err
    open(marker, 'w').close()


def flush(path=None):
    """Wait until the file(s) this process is writing are complete on disk (all of them when path is None).
    Must precede anything outside this process that reads them."""
    keys = [os.path.realpath(path)] if path is not None else list(dict.fromkeys(list(_WRITERS) + list(_FAILED)))
    for rp in keys:
        w = _WRITERS.get(rp)
        if w is not None:
            w.join()
        err = _FAILED.pop(rp, None)
        if err is not None:
            raise err


def _remember(rp, size, mtime_ns, t):
    """Keeps the sources it was asked to load plus the newest product; older products go."""
    if t is not None and rp.endswith('.bed') and not _is_source(rp):
        for k in [k for k in _RESIDENT if not _is_source(k) and k not in _WRITERS]:
            del _RESIDENT[k]
    _RESIDENT.pop(rp, None)
    _RESIDENT[rp] = (size, mtime_ns, t)
    src = [k for k in _RESIDENT if _is_source(k)]
    for k in src[:-2]:
        del _RESIDENT[k]


def _is_source(rp):
    return os.path.basename(rp) == 'alignment_iteration_1.bed' or not os.path.basename(rp).startswith('alignment_iteration_')


def reproject(eng, bed, refactor, scaffold_of, offset, length, reversed_, scaf_names):
    """bed: uint8 CUDA tensor with the whole iteration-1 BED.  The four arrays are indexed by
    the Engine's contig id (eng.names order); scaf_names: list[bytes].  Returns the new BED
    text as a uint8 CUDA tensor.  Raises DataError where the reference's loop would raise."""
    L, dev = eng.L, eng.device
    st = torch.cuda.current_stream().cuda_stream
    so = torch.from_numpy(np.ascontiguousarray(scaffold_of, np.int32)).to(dev)
    of = torch.from_numpy(np.ascontiguousarray(offset, np.int64)).to(dev)
    ln = torch.from_numpy(np.ascontiguousarray(length, np.int64)).to(dev)
    rv = torch.from_numpy(np.ascontiguousarray(reversed_, np.uint8)).to(dev)
    blob, noff = _blob(scaf_names)
    sb = torch.from_numpy(np.frombuffer(blob + b'\0', dtype=np.uint8).copy()).to(dev)
    sf = torch.from_numpy(noff.astype(np.uint32)).to(dev)
    count = torch.zeros(1, dtype=torch.int64, device=dev)
    cap = max(1024, bed.numel() // 64)
    while True:
        cand = torch.empty(cap, dtype=torch.int64, device=dev)
        check(L.salsa_ingest_stats_reset(eng.stats.data_ptr(), st))
        check(L.salsa_reproject_candidates(eng.ctx, bed.data_ptr(), bed.numel(), int(bool(refactor)), so.data_ptr(),
                                           cand.data_ptr(), cap, count.data_ptr(), eng.stats.data_ptr(), st))
        eng.stats_read()
        n = int(count.item())
        if n <= cap:
            break
        cap = n
    if n == 0:
        eng.last_candidate = 0
        return torch.empty(0, dtype=torch.uint8, device=dev)
    srt = torch.empty(n, dtype=torch.int64, device=dev)
    tb = C.c_size_t(0)
    check(L.salsa_sort_u64(cand.data_ptr(), srt.data_ptr(), n, bed.numel(), None, C.byref(tb), st))
    tmp = torch.empty(tb.value + 16, dtype=torch.uint8, device=dev)
    check(L.salsa_sort_u64(cand.data_ptr(), srt.data_ptr(), n, bed.numel(), tmp.data_ptr(), C.byref(tb), st))
    del cand, tmp
    nblk = L.salsa_reproject_blocks(n)
    lens = torch.empty(n, dtype=torch.int32, device=dev)
    blk = torch.empty(nblk + 1, dtype=torch.int64, device=dev)
    args = (eng.ctx, bed.data_ptr(), bed.numel(), int(bool(refactor)), srt.data_ptr(), n, so.data_ptr(), of.data_ptr(),
            ln.data_ptr(), rv.data_ptr(), sb.data_ptr(), sf.data_ptr(), lens.data_ptr(), blk.data_ptr())
    check(L.salsa_reproject_emit(*args, None, 0, eng.stats.data_ptr(), st))
    check(L.salsa_exclusive_scan_u64(blk.data_ptr(), nblk + 1, st))
    total = int(blk[nblk].item())
    eng.last_candidate = int(srt[n - 1].item())       # refactor variant: offset of the last run start (chunked passes carry it)
    eng.stats_read()
    out = torch.empty(total, dtype=torch.uint8, device=dev)
    if total:
        check(L.salsa_reproject_emit(*args, out.data_ptr(), total, eng.stats.data_ptr(), st))
    return out


def _chunked_pass(eng, src, dst, refactor, tables, chunk_bytes):
    """The same pass for a BED that does not fit the GPU beside its product: line-aligned chunks streamed from
    the file, the product appended to dst.  A chunk starts with the line that the sliding window still needs
    from before it: the previous chunk's last line (layout variant) or its last run start (refactor variant:
    the reference's prev_line only ever holds a run start, refactor_breaks.py:150-151,195)."""
    size = os.path.getsize(src)
    rp = os.path.realpath(dst)
    flush(dst)
    _RESIDENT.pop(rp, None)
    carry = b''
    with open(src, 'rb', buffering=0) as f, open(rp + '.part', 'wb') as out:
        pos = 0
        while pos < size:
            blk = f.read(min(chunk_bytes, size - pos))
            if pos + len(blk) < size:
                cut = blk.rfind(b'\n') + 1
                if cut == 0:
                    raise ValueError('BED line longer than the re-projection chunk near byte %d' % pos)
                blk = blk[:cut]
                f.seek(pos + cut)
            pos += len(blk)
            data = carry + blk
            t = torch.frombuffer(bytearray(data), dtype=torch.uint8).to(eng.device)
            res = reproject(eng, t, refactor, *tables)
            res.cpu().numpy().tofile(out)
            if refactor:
                a = eng.last_candidate                      # 0: the carried line is still the open run's start
            else:
                a = data.rfind(b'\n', 0, len(data) - 1) + 1 if data.endswith(b'\n') else data.rfind(b'\n') + 1
            e = data.find(b'\n', a)
            carry = data[a:e + 1] if e >= 0 else data[a:] + b'\n'
            del t, res
    os.replace(rp + '.part', rp)


def clean_pass(eng):
    """-> bed_pass(src, dst, names, cut, flush_every) for salsa_b200.host.clean.clean: the BED rewrite of correct.py:62-92
    on the device (salsa_clean_bed_plan / salsa_clean_bed_emit).  Line starts and the prefix sums between the two
    kernels are torch ops; the reference's lost tail (its last, partial buffer of `flush_every` lines is never written)
    is applied to the ranks of the kept lines.  The cleaned BED stays resident for the steps that follow."""
    from ._lib import DataError

    def run(src, dst, names, cut, flush_every):
        dev = eng.device
        bed = resident_bed(src, dev, load=True)
        nb = [n.encode('latin-1') for n in names]
        eng.set_contigs(nb, np.zeros(len(nb), np.int64))
        idx = {n: i for i, n in enumerate(eng.names)}
        cut_arr = np.full(max(len(eng.names), 1), -1, np.int64)
        for k, v in cut.items():
            j = idx.get(k.encode('latin-1'))
            if j is not None:
                if v < 0:
                    raise ValueError('negative break position for %r' % k)
                cut_arr[j] = v
        d_cut = torch.from_numpy(cut_arr).to(dev)
        n = bed.numel()
        st = torch.cuda.current_stream().cuda_stream
        starts = torch.nonzero(bed == 10).flatten() + 1
        starts = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), starts]) if n else starts
        if starts.numel() and int(starts[-1]) >= n:
            starts = starts[:-1]                                   # the file ends with a newline: no line starts there
        nl = starts.numel()
        out_len = torch.empty(max(nl, 1), dtype=torch.int32, device=dev)
        err = torch.empty(1, dtype=torch.int64, device=dev)
        check(eng.L.salsa_clean_bed_plan(eng.ctx, bed.data_ptr(), n, d_cut.data_ptr(), starts.data_ptr(), nl, out_len.data_ptr(),
                                         err.data_ptr(), st))
        e = int(err.item())
        if e != -1:
            raise DataError(e & 0xff, (e >> 8) & ((1 << 56) - 1))
        out_len = out_len[:nl]
        keep = out_len > 0
        rank = torch.cumsum(keep, 0)                               # 1-based rank among the kept lines
        kept = int(rank[-1]) if nl else 0
        limit = kept // flush_every * flush_every                  # correct.py:87-92: the last partial buffer is lost
        out_len = torch.where(keep & (rank <= limit), out_len, torch.zeros_like(out_len))
        off = torch.cumsum(out_len.to(torch.int64), 0) - out_len
        total = int((off[-1] + out_len[-1]).item()) if nl else 0
        out = torch.empty(total, dtype=torch.uint8, device=dev)
        if nl and total:
            check(eng.L.salsa_clean_bed_emit(eng.ctx, bed.data_ptr(), n, d_cut.data_ptr(), starts.data_ptr(), nl, out_len.data_ptr(),
                                             off.data_ptr(), out.data_ptr(), st))
        torch.cuda.current_stream().synchronize()
        with open(dst, 'wb') as f:
            pos = 0
            while pos < total:
                k = min(CHUNK, total - pos)
                out[pos:pos + k].cpu().numpy().tofile(f)
                pos += k
        st_ = os.stat(dst)
        _remember(os.path.realpath(dst), st_.st_size, st_.st_mtime_ns, out)
        # run_pipeline.py:125 moves the file to alignment_iteration_1.bed (same size, same mtime): known under that name too
        final = os.path.join(os.path.dirname(os.path.realpath(dst)), 'alignment_iteration_1.bed')
        if final != os.path.realpath(dst):
            _remember(final, st_.st_size, st_.st_mtime_ns, out)
    return run


def re_pass(eng):
    """-> re_pass(scaffolds, contig_length, re_counts) for salsa_b200.host.reproject.update_bed: the per-scaffold RE
    proration of layout_unitigs.py:736-814 / refactor_breaks.py:63-131 on the device (salsa_scaffold_re)."""
    def run(scaffolds, contig_length, re_counts):
        keys = list(scaffolds.keys())
        cid, clen, rl, rr = {}, [], [], []
        off, flat, first_b = [0], [], []
        for key in keys:
            path = scaffolds[key]
            for i in range(0, len(path), 2):
                contig = path[i].split(':')[0]
                j = cid.get(contig)
                if j is None:
                    j = cid[contig] = len(clen)
                    clen.append(int(contig_length[contig]))
                    left, right = re_counts[contig]                       # KeyError like the reference's lookup
                    rl.append(left)
                    rr.append(right)
                flat.append(j)
            off.append(len(flat))
            first_b.append(1 if path[0].split(':')[1] == 'B' else 0)
        n = len(keys)
        if n == 0:
            return {}
        dev = eng.device
        t = lambda a, dt: torch.tensor(a, dtype=dt, device=dev)           # noqa: E731
        d_off, d_flat, d_fb = t(off, torch.int64), t(flat, torch.int32), t(first_b, torch.uint8)
        d_len, d_rl, d_rr = t(clen, torch.int64), t(rl, torch.int64), t(rr, torch.int64)
        ol, orr = torch.empty(n, dtype=torch.int64, device=dev), torch.empty(n, dtype=torch.int64, device=dev)
        check(eng.L.salsa_scaffold_re(d_off.data_ptr(), d_flat.data_ptr(), d_fb.data_ptr(), d_len.data_ptr(), d_rl.data_ptr(),
                                      d_rr.data_ptr(), n, ol.data_ptr(), orr.data_ptr(), torch.cuda.current_stream().cuda_stream))
        ol, orr = ol.cpu().tolist(), orr.cpu().tolist()
        return {k: (a, b) for k, a, b in zip(keys, ol, orr)}
    return run


def bed_pass(eng, background=False, chunk_bytes=None):
    """-> bed_pass(src, dst, c2s, info, contig_length, refactor) for salsa_b200.host.reproject.update_bed.
    background: return while dst is still being written (the caller promises to flush() before anything
    outside this process reads it); the device copy serves every reader inside the process meanwhile.
    chunk_bytes: force the chunked pass (default: only when the BED does not fit the GPU three times over)."""
    def run(src, dst, c2s, info, contig_length, refactor):
        names = [k.encode('latin-1') for k in contig_length]
        idx = eng.set_contigs(names, np.array([contig_length[k] for k in contig_length], dtype=np.int64))
        n = eng.n
        scaf_keys = list(dict.fromkeys(c2s.values()))
        sidx = {k: i for i, k in enumerate(scaf_keys)}
        scaffold_of = np.full(n, -1, dtype=np.int32)
        offset = np.zeros(n, dtype=np.int64)
        length = np.zeros(n, dtype=np.int64)
        rev = np.zeros(n, dtype=np.uint8)
        for k, ln in contig_length.items():
            length[idx[k.encode('latin-1')]] = ln
        for contig, key in c2s.items():
            i = idx[contig.encode('latin-1')]          # KeyError: a placed contig without a length, as in the reference
            scaffold_of[i] = sidx[key]
            offset[i] = info[contig][0]
            rev[i] = info[contig][2] != 'FOW'
        tables = (scaffold_of, offset, length, rev, [k.encode('latin-1') for k in scaf_keys])
        free, _ = torch.cuda.mem_get_info(eng.device)
        if chunk_bytes is not None or (resident_bed(src, eng.device, load=False) is None and os.path.getsize(src) * 3 > free):
            _chunked_pass(eng, src, dst, refactor, tables, chunk_bytes or (4 << 30))
            return
        bed = resident_bed(src, eng.device)
        out = reproject(eng, bed, refactor, *tables)
        write_behind(dst, out)
        if not background:
            flush(dst)
    run.after_write = after_write
    return run
