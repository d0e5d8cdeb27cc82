"""GPU BED pass of update_bed: alignment_iteration_1.bed -> alignment_iteration_N.bed in
scaffold coordinates (layout_unitigs.py:816-886, refactor_breaks.py:133-203), on top of
salsa_reproject_candidates / salsa_sort_u64 / salsa_reproject_emit.

The iteration-1 BED is read from disk once per process and then stays in HBM (every
iteration re-projects it twice); the BED each pass produces also stays resident, so the
make_links.py step that follows ingests it without touching the file again.
"""
import ctypes as C
import os

import numpy as np
import torch

from ._lib import check
from .engine import _blob

_RESIDENT = {}          # realpath -> (size, mtime_ns, uint8 CUDA tensor)


def resident_bed(path, device, load=True):
    """uint8 CUDA tensor holding the file, from the cache when its size and mtime still match."""
    rp = os.path.realpath(path)
    st = os.stat(rp)
    hit = _RESIDENT.get(rp)
    if hit and hit[0] == st.st_size and hit[1] == st.st_mtime_ns and hit[2].device == device:
        return hit[2]
    if not load:
        return None
    free, _ = torch.cuda.mem_get_info(device)
    if st.st_size * 3 > free:
        raise MemoryError('%s (%d bytes) does not fit the GPU beside its re-projection; chunked re-projection is not '
                          'implemented' % (path, st.st_size))
    if st.st_size == 0:
        t = torch.empty(0, dtype=torch.uint8, device=device)
    else:
        t = torch.from_numpy(np.fromfile(rp, dtype=np.uint8)).to(device)
    _remember(rp, st.st_size, st.st_mtime_ns, t)
    return t


def _remember(rp, size, mtime_ns, t):
    """Keeps the sources it was asked to load plus the newest product; older products go."""
    if t is not None and rp.endswith('.bed') and not _is_source(rp):
        for k in [k for k in _RESIDENT if not _is_source(k)]:
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
        return torch.empty(0, dtype=torch.uint8, device=dev)
    srt = torch.empty(n, dtype=torch.int64, device=dev)
    tb = C.c_size_t(0)
    check(L.salsa_sort_u64(cand.data_ptr(), srt.data_ptr(), n, None, C.byref(tb), st))
    tmp = torch.empty(tb.value + 16, dtype=torch.uint8, device=dev)
    check(L.salsa_sort_u64(cand.data_ptr(), srt.data_ptr(), n, tmp.data_ptr(), C.byref(tb), st))
    del cand, tmp
    off = torch.zeros(n + 1, dtype=torch.int64, device=dev)
    args = (eng.ctx, bed.data_ptr(), bed.numel(), int(bool(refactor)), srt.data_ptr(), n, so.data_ptr(), of.data_ptr(),
            ln.data_ptr(), rv.data_ptr(), sb.data_ptr(), sf.data_ptr(), off.data_ptr())
    check(L.salsa_reproject_emit(*args, None, 0, eng.stats.data_ptr(), st))
    check(L.salsa_exclusive_scan_u64(off.data_ptr(), n + 1, st))
    total = int(off[n].item())
    eng.stats_read()
    out = torch.empty(total, dtype=torch.uint8, device=dev)
    if total:
        check(L.salsa_reproject_emit(*args, out.data_ptr(), total, eng.stats.data_ptr(), st))
    return out


def bed_pass(eng):
    """-> bed_pass(src, dst, c2s, info, contig_length, refactor) for salsa_b200.host.reproject.update_bed."""
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
        bed = resident_bed(src, eng.device)
        out = reproject(eng, bed, refactor, scaffold_of, offset, length, rev, [k.encode('latin-1') for k in scaf_keys])
        host = out.cpu().numpy()
        with open(dst, 'wb') as f:
            host.tofile(f)
        st = os.stat(dst)
        _remember(os.path.realpath(dst), st.st_size, st.st_mtime_ns, out)
    return run
