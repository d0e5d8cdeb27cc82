"""Host side of restriction-site counting (RE_sites.py): FASTA record bookkeeping
around the device kernels of salsa_b200/csrc/re_sites.cu.

The host touches only the header lines (names) and the enzyme string; every
pass over sequence bytes -- record boundaries, sequence lengths, packing and
pattern counting -- is a kernel.
"""
import ctypes as C
import re

import numpy as np
import torch

from ._lib import check


class FastaError(Exception):
    """A condition that makes the reference exit 1 or raise (RE_sites.py:13-15,19,22,48-50)."""


def expand_enzymes(enzyme):
    """RE_sites.py:45-58: comma list, 'N' -> the same base in every N position, in G,A,T,C order."""
    out = []
    for each in enzyme.replace(' ', '').split(','):
        if not re.match("^[ACGTN]+$", each):
            raise FastaError("Error, enzyme should be restriction site sequence (e.g. AACTT) not enzyme name or DNASE, "
                             "you input %s" % each)
        if 'N' in each:
            for b in 'GATC':
                out.append(each.replace('N', b))
        else:
            out.append(each)
    return out


class _DeviceView:
    """Byte access to a FASTA that lives only in device memory (bench / pipelines that keep it in HBM)."""

    def __init__(self, t):
        self.t = t
        self.size = t.numel()

    def __getitem__(self, key):
        return self.t[key].cpu().numpy()


def count_sites(eng, fasta, enzyme, timers=None, device_copy=None):
    """fasta: host uint8 numpy array (np.memmap works) or a uint8 CUDA tensor.  Returns
    (names, left, right, seqlen) in first-appearance order of record names; a repeated
    name keeps its first position and its last sequence (dict assignment, RE_sites.py:19).
    timers: optional dict that receives (start, end) CUDA event pairs per stage.
    device_copy: the same bytes already on the device (the header lines are still read from `fasta` on the host)."""
    L, dev = eng.L, eng.device
    st = torch.cuda.current_stream().cuda_stream
    d = device_copy
    if isinstance(fasta, torch.Tensor):
        d = fasta
        fasta = _DeviceView(d)
    n = int(fasta.size)
    if n == 0:
        return [], np.zeros(0, np.int64), np.zeros(0, np.int64), np.zeros(0, np.int64)
    if fasta[0:1][0] != ord('>'):
        raise FastaError('sequence line before the first header (the reference raises KeyError, RE_sites.py:22)')
    patterns = None if enzyme == 'DNASE' else expand_enzymes(enzyme)          # validate before any GPU work
    if d is None:
        if n > (64 << 20):
            # a large host array (np.memmap of the assembly, or a rank's byte range of it): through pinned buffers, four
            # copying threads side by side, instead of a private copy followed by a pageable transfer
            from .reproject import _stream_in
            d = _stream_in(None, n, dev, src=np.ascontiguousarray(fasta) if not fasta.flags.c_contiguous else fasta)
        else:
            d = torch.from_numpy(np.array(fasta, copy=True) if not fasta.flags.writeable else np.ascontiguousarray(fasta)).to(dev)

    def tick(name):
        if timers is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            timers.setdefault(name, []).append(ev)

    tick('headers')
    cap = 1 << 16
    while True:
        pos = torch.empty(cap, dtype=torch.int64, device=dev)
        cnt = torch.zeros(1, dtype=torch.int32, device=dev)
        check(L.salsa_fasta_headers(d.data_ptr(), n, pos.data_ptr(), cnt.data_ptr(), cap, st))
        k = int(cnt.item())
        if k <= cap:
            break
        cap = k
    tick('headers')
    hp = np.sort(pos[:k].cpu().numpy())
    names, sbeg, send = [], np.zeros(k, np.uint64), np.zeros(k, np.uint64)
    for i, p in enumerate(hp):
        p = int(p)
        w = 256
        while True:
            seg = fasta[p:min(p + w, n)]
            nl = np.flatnonzero(seg == 10)
            if len(nl) or p + w >= n:
                break
            w *= 4
        he = p + int(nl[0]) if len(nl) else n
        ln = bytes(fasta[p:he])
        if b':' in ln:                                                     # RE_sites.py:13-15
            raise FastaError("Error, your fasta records contain a ':' (%s). SALSA requires fasta to not contain this "
                             "character, please update your fasta and re-try" % ln[1:].rstrip().decode('latin-1'))
        toks = ln[1:].rstrip().split()
        if not toks:
            raise FastaError('empty fasta header (IndexError in the reference, RE_sites.py:19)')
        names.append(toks[0])
        sbeg[i] = min(he + 1, n)
        send[i] = int(hp[i + 1]) if i + 1 < k else n
    dsb = torch.from_numpy(sbeg.view(np.int64)).to(dev)
    dse = torch.from_numpy(send.view(np.int64)).to(dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    nt = L.salsa_fasta_tiles(n)
    tiles = torch.zeros(nt + 1, dtype=torch.int64, device=dev)
    tick('count')
    check(L.salsa_fasta_count(d.data_ptr(), n, dsb.data_ptr(), dse.data_ptr(), k, tiles.data_ptr(), err.data_ptr(), st))
    check(L.salsa_exclusive_scan_u64(tiles.data_ptr(), nt, st))
    pbeg = torch.zeros(k, dtype=torch.int64, device=dev)
    plen = torch.zeros(k, dtype=torch.int64, device=dev)
    check(L.salsa_fasta_span_lengths(d.data_ptr(), n, dsb.data_ptr(), dse.data_ptr(), k, tiles.data_ptr(), pbeg.data_ptr(),
                                     plen.data_ptr(), err.data_ptr(), st))
    tick('count')
    seqlen = plen.cpu().numpy()
    if int(err.item()):
        raise FastaError('a run of trailing whitespace longer than 1024 bytes (library limit)')
    if patterns is None:                                                   # DNASE, RE_sites.py:39-43
        left = seqlen // 2
        right = seqlen // 2
    else:
        blk = np.zeros(k + 1, dtype=np.int64)
        blk[1:] = np.cumsum((seqlen + 63) // 64)
        nblk = int(blk[-1])
        dblk = torch.from_numpy(blk).to(dev)
        planes = torch.zeros(max(3 * nblk, 1), dtype=torch.int64, device=dev)
        tick('pack')
        check(L.salsa_fasta_pack(d.data_ptr(), n, dsb.data_ptr(), dse.data_ptr(), k, tiles.data_ptr(), pbeg.data_ptr(),
                                 dblk.data_ptr(), planes.data_ptr(), err.data_ptr(), st))
        tick('pack')
        dl = torch.zeros(k, dtype=torch.int64, device=dev)
        dr = torch.zeros(k, dtype=torch.int64, device=dev)
        nborder = sum(1 for p in patterns if any(p[:j] == p[-j:] for j in range(1, len(p))))
        occ = torch.zeros(max(nborder * nblk, 1), dtype=torch.int64, device=dev) if nborder else None
        blob = b''.join(p.encode() + b'\0' for p in patterns)
        tick('re_count')
        check(L.salsa_re_count(planes.data_ptr(), dblk.data_ptr(), plen.data_ptr(), k, nblk, blob, len(patterns), dl.data_ptr(),
                               dr.data_ptr(), occ.data_ptr() if occ is not None else None, st))
        tick('re_count')
        left, right = dl.cpu().numpy(), dr.cpu().numpy()
        if timers is not None:
            timers['_sizes'] = dict(fasta_bytes=n, bases=int(seqlen.sum()), plane_bytes=int(nblk) * 24, npatterns=len(patterns))
    # dict semantics: first position, last value
    order, last = [], {}
    for i, nm in enumerate(names):
        if nm not in last:
            order.append(nm)
        last[nm] = i
    idx = np.array([last[nm] for nm in order], dtype=np.int64)
    return order, left[idx], right[idx], seqlen[idx]
