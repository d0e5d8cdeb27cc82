"""Python face of the coverage kernels (salsa_cov_scan_*, salsa_cov_analyse).

The drop-in programs for break_contigs_start / break_contigs are the C++ hosts in
salsa_b200/csrc/host (the reference's tools are C++); this module drives the same
C-ABI entry points from Python for benchmarks and tests.
"""
import ctypes as C

import numpy as np
import torch

from ._lib import check

CHUNK = 4096
NOSEED = 2 ** 64 - 1


class Seg(C.Structure):
    _fields_ = [('src_off', C.c_uint64), ('dst_off', C.c_uint64), ('seed_idx', C.c_uint64), ('len', C.c_int64),
                ('lo', C.c_int64), ('hi', C.c_int64)]


def layout(len_by_id):
    """Element offsets of the per-contig difference arrays (len + 10 ints, chunk aligned) for contigs
    >= 1 Mb (break_contigs_start.cpp:325,335); UINT64_MAX elsewhere.  Returns (covoff, total)."""
    covoff = np.full(len(len_by_id), NOSEED, dtype=np.uint64)
    tot = 0
    for i, ln in enumerate(len_by_id):
        if ln >= 1_000_000:
            covoff[i] = tot
            tot += (int(ln) + 10 + CHUNK - 1) // CHUNK * CHUNK
    return covoff, tot


def scan(eng, bed, len_by_id, covoff, ndiff, joins=None, first_chunk=True, diff=None, touched=None, base_offset=0):
    """BED text (uint8 CUDA tensor) -> difference arrays.  joins: (join_off uint64[n+1], join_pos int64[]) of
    the KEPT join positions per contig id (ascending) selects the break_contigs variant."""
    L, dev = eng.L, eng.device
    st = torch.cuda.current_stream().cuda_stream
    n = len(len_by_id)
    d_off = torch.from_numpy(covoff.view(np.int64)).to(dev)
    d_len = torch.from_numpy(np.ascontiguousarray(len_by_id, np.int64)).to(dev)
    if diff is None:
        diff = torch.zeros(ndiff + 1, dtype=torch.int32, device=dev)
    if touched is None:
        touched = torch.zeros(n + 1, dtype=torch.uint8, device=dev)
    if joins is None:
        check(L.salsa_cov_scan_start(eng.ctx, bed.data_ptr(), bed.numel(), base_offset, int(first_chunk), d_off.data_ptr(),
                                     d_len.data_ptr(), diff.data_ptr(), touched.data_ptr(), eng.stats.data_ptr(), st))
    else:
        joff = torch.from_numpy(np.ascontiguousarray(joins[0]).view(np.int64)).to(dev)
        jpos = torch.from_numpy(np.ascontiguousarray(np.append(joins[1], 0), np.int64)).to(dev)
        check(L.salsa_cov_scan_joins(eng.ctx, bed.data_ptr(), bed.numel(), base_offset, int(first_chunk), d_off.data_ptr(),
                                     d_len.data_ptr(), joff.data_ptr(), jpos.data_ptr(), diff.data_ptr(), touched.data_ptr(),
                                     eng.stats.data_ptr(), st))
    return diff, touched


def analyse(eng, diff, cov, segs):
    """Prefix sums + the 11 Kadane windows per segment.  Returns (sum[n], start[n,11], end[n,11])."""
    L = eng.L
    n = len(segs)
    arr = (Seg * n)(*segs)
    sb = L.salsa_cov_scratch_bytes(arr, n)
    scratch = torch.empty(sb // 8 + 1, dtype=torch.int64, device=eng.device)
    hs = np.zeros(n, np.int64)
    st = np.zeros(n * 11, np.int64)
    en = np.zeros(n * 11, np.int64)
    check(L.salsa_cov_analyse(diff.data_ptr(), cov.data_ptr(), arr, n, scratch.data_ptr(), sb, hs.ctypes.data, st.ctypes.data,
                              en.ctypes.data, torch.cuda.current_stream().cuda_stream))
    return hs, st.reshape(n, 11), en.reshape(n, 11)
