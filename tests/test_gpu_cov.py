"""GPU parity tests of the coverage path (break_contigs_start / break_contigs):
the drop-in host programs against the compiled reference's golden outputs, and the
prefix/Kadane kernels against the oracle on many small segments (tie rules)."""
import ctypes as C
import json
import os
import subprocess
import tempfile

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, 'salsa_b200', 'bin')
GOLD = os.path.join(ROOT, 'tests', 'golden')


class Seg(C.Structure):
    _fields_ = [('src_off', C.c_uint64), ('dst_off', C.c_uint64), ('seed_idx', C.c_uint64), ('len', C.c_int64),
                ('lo', C.c_int64), ('hi', C.c_int64)]


def analyse(L, torch, diff, segs, inplace):
    dev = torch.device('cuda', 0)
    d = torch.from_numpy(np.ascontiguousarray(diff, np.int32)).to(dev)
    n = len(segs)
    arr = (Seg * n)(*segs)
    cov = d if inplace else torch.zeros(max(sum(s.len for s in segs), 1), dtype=torch.int32, device=dev)
    sb = L.salsa_cov_scratch_bytes(arr, n)
    scratch = torch.zeros(sb // 8 + 1, dtype=torch.int64, device=dev)
    hs = np.zeros(n, np.int64); st = np.zeros(n * 11, np.int64); en = np.zeros(n * 11, np.int64)
    rc = L.salsa_cov_analyse(d.data_ptr(), cov.data_ptr(), arr, n, scratch.data_ptr(), sb, hs.ctypes.data, st.ctypes.data,
                             en.ctypes.data, None)
    assert rc == 0, L.salsa_last_error()
    return hs, st.reshape(n, 11), en.reshape(n, 11), cov.cpu().numpy()


@pytest.mark.parametrize('seed', [0, 1, 2])
def test_prefix_and_kadane_small_segments(oracle_lib, seed):
    """Many short segments with few distinct coverage values: ties everywhere."""
    import torch
    from salsa_b200 import _lib
    L = _lib.lib()
    rng = np.random.default_rng(seed)
    lens = [int(x) for x in rng.choice([1, 2, 3, 5, 17, 64, 255, 256, 257, 4095, 4096, 4097, 9000, 30000], 40)]
    segs, diffs, off = [], [], 0
    for ln in lens:
        kind = rng.integers(0, 3)
        if kind == 0:
            d = rng.integers(-1, 2, ln)
        elif kind == 1:
            d = np.zeros(ln, np.int64)
            for _ in range(max(1, ln // 50)):
                a, b = sorted(rng.integers(0, ln, 2))
                d[a] += 1
                d[b] -= 1
        else:
            d = rng.integers(-3, 4, ln) * (rng.random(ln) < 0.05)
        lo = int(rng.integers(0, max(ln // 3, 1)))
        hi = int(rng.integers(lo + 1, ln + 1))
        segs.append(Seg(off, off, 2 ** 64 - 1, ln, lo, hi))
        diffs.append(d.astype(np.int32))
        off += (ln + 4095) // 4096 * 4096
    flat = np.zeros(off, np.int32)
    for s, d in zip(segs, diffs):
        flat[s.src_off:s.src_off + s.len] = d
    hs, st, en, cov = analyse(L, torch, flat, segs, inplace=True)
    for k, (s, d) in enumerate(zip(segs, diffs)):
        osum, ost, oen = oracle_lib.cov_analyse(d, 0, s.lo, s.hi)
        assert hs[k] == osum
        assert (cov[s.src_off:s.src_off + s.len] == np.cumsum(d)).all()
        assert (st[k] == ost).all() and (en[k] == oen).all(), (k, s.len, s.lo, s.hi, st[k], ost, en[k], oen)


def test_overlapping_seeded_windows(oracle_lib):
    """break_contigs.cpp-style windows: overlapping sources, seeded by the raw diff[start-1], separate output."""
    import torch
    from salsa_b200 import _lib
    L = _lib.lib()
    rng = np.random.default_rng(9)
    n = 60000
    flat = (rng.integers(-2, 3, n) * (rng.random(n) < 0.1)).astype(np.int32)
    segs, dst = [], 0
    for start in (1, 5000, 7000, 20000, 31000):
        ln = 25000
        segs.append(Seg(start, dst, start - 1, ln, ln // 3, 2 * ln // 3))
        dst += ln
    hs, st, en, cov = analyse(L, torch, flat, segs, inplace=False)
    for k, s in enumerate(segs):
        d = flat[s.src_off:s.src_off + s.len]
        osum, ost, oen = oracle_lib.cov_analyse(d, int(flat[s.seed_idx]), s.lo, s.hi)
        assert hs[k] == osum and (st[k] == ost).all() and (en[k] == oen).all()
        assert (cov[s.dst_off:s.dst_off + s.len] == flat[s.seed_idx] + np.cumsum(d)).all()


def _run(prog, args):
    r = subprocess.run([os.path.join(BIN, prog)] + args, capture_output=True)
    assert r.returncode == 0, r.stderr.decode()
    return sorted(r.stdout.decode().splitlines()), r.stderr.decode()


@pytest.mark.parametrize('case', sorted(cases.COV_START_CASES))
def test_break_contigs_start_golden(case):
    g = json.load(open(os.path.join(GOLD, 'coverage_ref.json')))
    seed, lens, holes, npairs = cases.COV_START_CASES[case]
    names, L, bed = cases.coverage_case(seed, lens, holes, npairs)
    d = tempfile.mkdtemp()
    open(d + '/b.bed', 'wb').write(bed)
    open(d + '/l.txt', 'wb').write(cases.lengths_text(names, L))
    out, err = _run('break_contigs_start', ['-a', d + '/b.bed', '-l', d + '/l.txt', '-s', '100'])
    assert out == g['start'][case]             # stdout of the UNMODIFIED reference binary
    assert 'bedfile loaded' in err
    out2, _ = _run('break_contigs_start', ['--alignment=' + d + '/b.bed', '--contiglen', d + '/l.txt', '--min_size=100'])
    assert out2 == out


@pytest.mark.parametrize('case', sorted(cases.COV_JOIN_CASES))
def test_break_contigs_golden(case):
    g = json.load(open(os.path.join(GOLD, 'coverage_ref.json')))
    seed, lens, holes, npairs, bp = cases.COV_JOIN_CASES[case]
    names, L, bed = cases.coverage_case(seed, lens, holes, npairs, scaffold_style=True)
    bpn = {names[k]: v for k, v in bp.items()}
    d = tempfile.mkdtemp()
    open(d + '/b.bed', 'wb').write(bed)
    open(d + '/l.txt', 'wb').write(cases.lengths_text(names, L))
    open(d + '/bp.txt', 'wb').write(cases.breakpoints_text(bpn))
    out, _ = _run('break_contigs', ['-a', d + '/b.bed', '-b', d + '/bp.txt', '-l', d + '/l.txt', '-i', '2', '-s', '100'])
    assert out == g['joins'][case]


def test_against_live_reference_extra_cases(oracle_lib):
    """More shapes than the committed fixtures, checked live against the compiled
    reference when it travelled with the repo (oracle/_ref), else against the oracle."""
    ref_s, ref_j = oracle_lib.ref_binary('break_contigs_start'), oracle_lib.ref_binary('break_contigs')
    d = tempfile.mkdtemp()
    extra = [(31, [1_000_000, 1_000_001, 999_999], {0: [400_000], 1: [650_000]}, 150_000),
             (32, [3_000_000], {0: [1_000_000, 2_100_000]}, 200_000)]
    for seed, lens, holes, npairs in extra:
        names, L, bed = cases.coverage_case(seed, lens, holes, npairs, overlap_frac=0.2)
        # a malformed line two thirds in: the reference stops reading there
        cut = bed.rfind(b'\n', 0, len(bed) * 2 // 3) + 1
        bed2 = bed[:cut] + b'ctg0000\tnot_a_number\t5\tx/1\n' + bed[cut:]
        for tag, text in (('a', bed), ('b', bed2), ('c', b'\n\n' + bed)):
            open(d + '/b.bed', 'wb').write(text)
            open(d + '/l.txt', 'wb').write(cases.lengths_text(names, L))
            got, _ = _run('break_contigs_start', ['-a', d + '/b.bed', '-l', d + '/l.txt', '-s', '100'])
            if ref_s:
                want = sorted(subprocess.run([ref_s, '-a', d + '/b.bed', '-l', d + '/l.txt', '-s', '100'],
                                             capture_output=True, check=True).stdout.decode().splitlines())
            else:
                want = sorted('%s\t%d' % (names[c].decode(), p) for c, p in oracle_lib.cov_start(text, names, L))
            assert got == want, (seed, tag)
    seed, lens, holes, npairs, bp = (41, [9_000_000, 4_500_000, 1_500_000], {0: [2_500_000, 6_000_000], 1: [2_250_000]}, 400_000,
                                     {0: [2_500_000, 4_400_000, 6_000_000, 9_000_000], 1: [2_250_000, 4_500_000], 2: [700_000, 1_500_000]})
    names, L, bed = cases.coverage_case(seed, lens, holes, npairs, scaffold_style=True)
    bpn = {names[k]: v for k, v in bp.items()}
    open(d + '/b.bed', 'wb').write(bed)
    open(d + '/l.txt', 'wb').write(cases.lengths_text(names, L))
    open(d + '/bp.txt', 'wb').write(cases.breakpoints_text(bpn))
    got, _ = _run('break_contigs', ['-a', d + '/b.bed', '-b', d + '/bp.txt', '-l', d + '/l.txt', '-i', '3', '-s', '100'])
    if ref_j:
        want = sorted(subprocess.run([ref_j, '-a', d + '/b.bed', '-b', d + '/bp.txt', '-l', d + '/l.txt', '-i', '3', '-s', '100'],
                                     capture_output=True, check=True).stdout.decode().splitlines())
        assert got == want
    rows, tot, sus = oracle_lib.cov_joins(bed, names, L, bpn)
    assert 'Total Joins = %d' % tot in got and 'Suspicious Joins = %d' % sus in got
