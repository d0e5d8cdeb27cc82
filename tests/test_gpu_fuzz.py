"""Structure fuzz of the BED scanner against the oracle: field widths around every fast-path limit (contig
names around 16 and 63 bytes, read names around 8 and 62 bytes, digit runs around 12, lines up to the
1024-byte limit), every separator str.split() accepts, signs, several '/', extra columns, empty lines, no
final newline -- dense and sparse tables, whole buffer and chunked.  Whatever the oracle makes of a file
(a link table, or the error the reference would die with), the CUDA path must make the same of it."""
import numpy as np
import pytest

from test_gpu_ingest import eng, gpu_dict, oracle_dict  # noqa: F401

pytestmark = pytest.mark.gpu


def build(seed, nlines, error_free=True):
    rng = np.random.default_rng(seed)
    lens = [1, 2, 7, 8, 9, 15, 16, 17, 18, 24, 31, 32, 33, 47, 62, 63, 64, 70, 120]
    names = []
    for i, ln in enumerate(lens):
        for j in range(2):
            stem = ('c%d_%d_' % (i, j)).encode()
            names.append((stem + b'x' * max(0, ln - len(stem)))[:max(ln, len(stem))])
    names = sorted(set(names))
    L = [int(rng.integers(1000, 9000)) for _ in names]
    seps = [b'\t', b' ', b'\t', b'\t', b'  ', b'\t ', b'\x0b', b'\x0c', b'\r']
    rlens = [1, 2, 6, 7, 8, 9, 14, 30, 31, 32, 33, 40, 61, 62, 63, 64, 90]
    out, r = [], 0
    for i in range(nlines):
        if rng.random() < 0.55:
            r += 1
        c = names[int(rng.integers(len(names)))]
        if not error_free and rng.random() < 0.01:
            c = b'unknown_contig'
        nd = int(rng.choice([1, 3, 7, 11, 12, 13, 14]))
        s = int(rng.integers(0, 10 ** min(nd, 9)))
        e = s + int(rng.integers(1, 300))
        stxt = (b'+' if rng.random() < 0.03 else b'') + (b'%d' % s).rjust(nd, b'0')
        rl = int(rng.choice(rlens))
        rd = (b'r%d_' % r + b'y' * rl)[:max(rl, len(b'r%d_' % r))]
        u = rng.random()
        if u < 0.6:
            rd += b'/%d' % int(rng.integers(1, 3))
        elif u < 0.7:
            rd += b'/1/extra/2'
        sep = lambda: seps[int(rng.integers(len(seps)))]          # noqa: E731
        tail = [b'', b'\t60\t+', b' extra col', b'\r', b'\t' * 3, b'\t' + b'z' * int(rng.integers(0, 700))][int(rng.integers(6))]
        lead = b' ' if rng.random() < 0.02 else b''
        line = lead + c + sep() + stxt + sep() + b'%d' % e + sep() + rd + tail
        if rng.random() < 0.004:
            line = b''
        if not error_free and rng.random() < 0.002:
            line = c + b'\t12\t' + b'x9'
        out.append(line[:1000])
    return names, L, b'\n'.join(out) + (b'\n' if seed % 2 else b'')


def clean(bed):
    """Drop what would make the reference raise (empty lines) so that the comparison is about counts."""
    return b'\n'.join(ln for ln in bed.split(b'\n') if ln.strip()) + (b'\n' if bed.endswith(b'\n') else b'')


@pytest.mark.parametrize('mode', ['dense', 'sparse'])
@pytest.mark.parametrize('seed', list(range(12)))
def test_structures(eng, oracle_lib, seed, mode):
    names, L, bed = build(seed, 12000)
    bed = clean(bed)
    want, res = oracle_dict(oracle_lib, bed, names, L)
    got, st = gpu_dict(eng, bed, names, L, mode=mode, offset=seed % 16)
    assert got == want and len(want) > 100
    assert st.pairs == res['npairs_counted']


@pytest.mark.parametrize('seed', [100, 101, 102, 103])
def test_chunked(eng, oracle_lib, seed):
    names, L, bed = build(seed, 9000)
    bed = clean(bed)
    want, _ = oracle_dict(oracle_lib, bed, names, L)
    # chunks start at a line start and overlap by exactly one line
    starts = [0] + [i + 1 for i, b in enumerate(bed) if b == 10 and i + 1 < len(bed)]
    rng = np.random.default_rng(seed)
    cuts = sorted(set(int(x) for x in rng.integers(1, len(starts) - 1, 6)))
    chunks, lo = [], 0
    for k in cuts:
        chunks.append((starts[lo], starts[k + 1] if k + 1 < len(starts) else len(bed)))
        lo = k
    chunks.append((starts[lo], len(bed)))
    got, _ = gpu_dict(eng, bed, names, L, chunks=chunks)
    assert got == want


@pytest.mark.parametrize('refactor', [False, True])
@pytest.mark.parametrize('seed', [300, 301, 302, 303])
def test_reprojection_structures(eng, tmp_path, seed, refactor):
    """The same line structures through the BED re-projection (candidate scan on the scanner's fast paths,
    emit by the byte-wise parser) against the oracle's twin of update_bed's pass."""
    from oracle import py_twin
    from salsa_b200 import reproject
    names, L, bed = build(seed, 8000)
    # update_bed needs a '/' in every read token (IndexError otherwise): give the bare ones a mate suffix
    lines = []
    for ln in clean(bed).split(b'\n'):
        f = ln.split()
        if len(f) >= 4 and b'/' not in f[3]:
            ln = ln.replace(f[3], f[3] + b'/2', 1)
        lines.append(ln)
    bed = b'\n'.join(lines)
    rng = np.random.default_rng(seed)
    snames = [n.decode('latin-1') for n in names]
    clen = dict(zip(snames, L))
    c2s, info, offs = {}, {}, [0] * 5
    for nm in snames:
        if rng.random() < 0.15:
            continue
        k = int(rng.integers(5))
        c2s[nm] = 'scaffold_%d' % (k + 1)
        info[nm] = (offs[k], offs[k] + clen[nm], 'FOW' if rng.random() < 0.5 else 'REV')
        offs[k] += clen[nm]
    src = str(tmp_path / 'alignment_iteration_1.bed')
    open(src, 'wb').write(bed)
    py_twin.reproject_bed(src, str(tmp_path / 'want.bed'), c2s, info, clen, refactor)
    reproject.bed_pass(eng)(src, str(tmp_path / 'got.bed'), c2s, info, clen, refactor)
    want = open(str(tmp_path / 'want.bed'), 'rb').read()
    assert open(str(tmp_path / 'got.bed'), 'rb').read() == want and len(want) > 1000


@pytest.mark.parametrize('seed', [200, 201, 202, 203, 204, 205])
def test_same_error_or_same_table(eng, oracle_lib, seed):
    import oracle
    from salsa_b200.engine import DataError
    names, L, bed = build(seed, 6000, error_free=False)
    try:
        want, _ = oracle_dict(oracle_lib, bed, names, L)
    except oracle.OracleError:
        with pytest.raises(DataError):
            gpu_dict(eng, bed, names, L)
        return
    got, _ = gpu_dict(eng, bed, names, L)
    assert got == want
