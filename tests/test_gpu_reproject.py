"""GPU parity tests of the BED re-projection (update_bed's pass over alignment_iteration_1.bed,
layout_unitigs.py:816-886 and refactor_breaks.py:133-203) against the oracle's twin: the output
file must be byte-identical, in both variants."""
import os
import sys

import numpy as np
import torch
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, 'salsa_b200', 'bin')


@pytest.fixture(scope='module')
def eng():
    sys.path.insert(0, BIN)
    from _common import engine
    return engine()


def placement(rng, names, lengths, nscaf, unplaced=0.2):
    """Random scaffolds over the contigs -> (c2s, info) as salsa_b200.host.reproject.placement builds them."""
    c2s, info = {}, {}
    offs = [0] * nscaf
    for nm, ln in zip(names, lengths):
        if rng.random() < unplaced:
            continue
        k = int(rng.integers(nscaf))
        key = 'scaffold_%d' % (k + 1)
        c2s[nm] = key
        info[nm] = (offs[k], offs[k] + ln, 'FOW' if rng.random() < 0.5 else 'REV')
        offs[k] += ln
    return c2s, info


def both(eng, tmp_path, bed, names, lengths, c2s, info, refactor):
    from oracle import py_twin
    from salsa_b200 import reproject
    src = str(tmp_path / 'src.bed')
    open(src, 'wb').write(bed)
    clen = dict(zip(names, lengths))
    py_twin.reproject_bed(src, str(tmp_path / 'want.bed'), c2s, info, clen, refactor)
    reproject.bed_pass(eng)(src, str(tmp_path / 'got.bed'), c2s, info, clen, refactor)
    return open(str(tmp_path / 'got.bed'), 'rb').read(), open(str(tmp_path / 'want.bed'), 'rb').read()


def random_bed(rng, names, lengths, npairs, same=0.15, noise=0.1):
    out = []
    n = len(names)
    for r in range(npairs):
        a = int(rng.integers(n))
        b = a if rng.random() < same else int(rng.integers(n))
        rows = [(a, '1'), (b, '2')]
        u = rng.random()
        if u < noise / 4:
            rows = rows[:1]                              # orphan mate
        elif u < noise / 2:
            rows = [(a, '2'), (b, '1')]                  # wrong mate order
        elif u < 3 * noise / 4:
            rows.append((int(rng.integers(n)), '2'))     # supplementary third line
        elif u < noise:
            rows = [(a, '1/x'), (b, '2/y')]              # more than one '/'
        for c, mate in rows:
            s = int(rng.integers(0, max(1, lengths[c] - 50)))
            sep = '\t' if rng.random() < 0.9 else '  '
            tail = '\t37\t+' if rng.random() < 0.5 else ''
            out.append('%s%s%d%s%d%sread%d/%s%s' % (names[c], sep, s, sep, s + 50, sep, r, mate, tail))
    return ('\n'.join(out) + '\n').encode('latin-1')


@pytest.mark.parametrize('refactor', [False, True])
@pytest.mark.parametrize('seed,ncontig,npairs', [(1, 5, 200), (2, 40, 5000), (3, 300, 60000)])
def test_random(eng, tmp_path, refactor, seed, ncontig, npairs):
    rng = np.random.default_rng(seed)
    names = ['ctg%d' % i if i % 7 else 'a_rather_long_contig_name_%d' % i for i in range(ncontig)]
    lengths = [int(x) for x in rng.integers(100, 100000, ncontig)]
    c2s, info = placement(rng, names, lengths, max(2, ncontig // 4))
    bed = random_bed(rng, names + ['not_in_the_assembly'], lengths + [5000], npairs)
    got, want = both(eng, tmp_path, bed, names, lengths, c2s, info, refactor)
    assert len(want) > 0
    assert got == want


@pytest.mark.parametrize('refactor', [False, True])
def test_known_answers(eng, tmp_path, refactor):
    names, lengths = ['A', 'B', 'C'], [100, 200, 300]
    c2s = {'A': 'scaffold_1', 'B': 'scaffold_1', 'C': 'scaffold_2'}
    info = {'A': (0, 100, 'FOW'), 'B': (100, 300, 'REV'), 'C': (0, 300, 'FOW')}
    bed = (b'A\t10\t20\tr1/1\n' b'C\t30\t40\tr1/2\n'         # kept by both: different scaffolds
           b'A\t1\t2\tr2/1\n' b'B\t50\t60\tr2/2\n'           # same scaffold: layout keeps, refactor drops
           b'B\t5\t6\tr3/1\n' b'B\t7\t8\tr3/2\n'             # same contig
           b'C\t9\t19\tr4/1\n' b'A\t3\t4\tr4/2')             # no final newline
    got, want = both(eng, tmp_path, bed, names, lengths, c2s, info, refactor)
    assert got == want
    if not refactor:
        assert got.startswith(b'scaffold_1\t10\t20\tr1/1\nscaffold_2\t30\t40\tr1/2\n'
                              b'scaffold_1\t1\t2\tr2/1\nscaffold_1\t240\t250\tr2/2\n')
    else:
        # r2 lands on one scaffold (dropped, :159); r3 is one contig; r4 closes against the stale previous line
        assert got == b'scaffold_1\t10\t20\tr1/1\nscaffold_2\t30\t40\tr1/2\nscaffold_2\t9\t19\tr4/1\nscaffold_1\t3\t4\tr4/2\n'


@pytest.mark.parametrize('refactor', [False, True])
def test_stale_previous_line(eng, tmp_path, refactor):
    """refactor variant: a same-contig line is skipped without becoming prev_line (refactor_breaks.py:150-151)."""
    names, lengths = ['A', 'B'], [100, 200]
    c2s = {'A': 'scaffold_1', 'B': 'scaffold_2'}
    info = {'A': (0, 100, 'FOW'), 'B': (0, 200, 'REV')}
    bed = b'A 1 2 x/1\nA 3 4 q/1\nA 5 6 q/1\nB 7 8 x/2\nB 9 10 z/2\nA 1 2 x/2\n'
    got, want = both(eng, tmp_path, bed, names, lengths, c2s, info, refactor)
    assert got == want
    if refactor:
        assert got == b'scaffold_1\t1\t2\tx/1\nscaffold_2\t192\t193\tx/2\n'
    else:
        assert got == b''


def test_tile_boundaries(eng, tmp_path):
    """Pairs on every side of the scanner's 32 KiB tiles, and a long run of same-contig lines."""
    names, lengths = ['A', 'B', 'C'], [10 ** 6, 10 ** 6, 10 ** 6]
    c2s = {'A': 'scaffold_1', 'B': 'scaffold_2'}
    info = {'A': (0, 10 ** 6, 'REV'), 'B': (5, 5 + 10 ** 6, 'FOW')}
    rows = []
    for r in range(20000):
        rows.append('A\t%d\t%d\tr%d/1' % (r, r + 30, r))
        rows.append('%s\t%d\t%d\tr%d/2' % ('B' if r % 3 else 'A', r * 2, r * 2 + 7, r))
        if r % 1000 == 0:
            rows += ['C\t1\t2\tpad%d/1' % r] * 40
    bed = ('\n'.join(rows) + '\n').encode()
    for refactor in (False, True):
        got, want = both(eng, tmp_path, bed, names, lengths, c2s, info, refactor)
        assert got == want and len(got) > 100000


@pytest.mark.parametrize('refactor', [False, True])
@pytest.mark.parametrize('bed', [b'A 1 2 r/1\nB 1 2 r\n', b'A 1 2 r/1\nB 1 2\n', b'A 1 x r/1\nB 1 2 r/2\n', b'A 1 2 r/1\n\nB 1 2 r/2\n'])
def test_errors_like_the_reference(eng, tmp_path, refactor, bed):
    """IndexError (missing field or '/') and ValueError (int()) of the reference -> DataError."""
    from salsa_b200.engine import DataError
    names, lengths = ['A', 'B'], [100, 200]
    c2s = {'A': 'scaffold_1', 'B': 'scaffold_2'}
    info = {'A': (0, 100, 'FOW'), 'B': (0, 200, 'FOW')}
    with pytest.raises((IndexError, ValueError)):
        from oracle import py_twin
        src = str(tmp_path / 's.bed')
        open(src, 'wb').write(bed)
        py_twin.reproject_bed(src, str(tmp_path / 'w.bed'), c2s, info, dict(zip(names, lengths)), refactor)
    from salsa_b200 import reproject
    with pytest.raises(DataError):
        reproject.bed_pass(eng)(src, str(tmp_path / 'g.bed'), c2s, info, dict(zip(names, lengths)), refactor)


def test_resident_product_feeds_make_links(eng, tmp_path):
    """The BED a pass leaves in HBM is what make_links.py ingests next (no re-read of the file)."""
    from salsa_b200 import reproject
    names, lengths = ['A', 'B'], [100, 200]
    c2s = {'A': 'scaffold_1', 'B': 'scaffold_2'}
    info = {'A': (0, 100, 'FOW'), 'B': (0, 200, 'FOW')}
    bed = b'A 1 2 r/1\nB 1 2 r/2\n'
    got, want = both(eng, tmp_path, bed, names, lengths, c2s, info, False)
    assert got == want
    t = reproject.resident_bed(str(tmp_path / 'got.bed'), eng.device, load=False)
    assert t is not None and bytes(t.cpu().numpy()) == got
    open(str(tmp_path / 'got.bed'), 'ab').write(b'X 1 2 q/1\n')          # changed on disk: cache entry must not be used
    assert reproject.resident_bed(str(tmp_path / 'got.bed'), eng.device, load=False) is None


def test_stream_in_with_several_readers(eng, tmp_path, monkeypatch):
    """A file of many chunks is read by several threads side by side (chunks round robin, own pinned buffers and
    streams): the device copy is the file, whatever the chunk size and the tail."""
    from salsa_b200 import reproject
    rng = np.random.default_rng(4)
    data = rng.integers(0, 256, 5_000_003, dtype=np.uint8).tobytes()
    p = str(tmp_path / 'big.bin')
    open(p, 'wb').write(data)
    for chunk in (1 << 16, 600_000, 1 << 23):
        monkeypatch.setattr(reproject, 'CHUNK', chunk)
        t = reproject._stream_in(p, len(data), eng.device)
        assert bytes(t.cpu().numpy()) == data, chunk
    open(p, 'wb').write(data[:1000])
    with pytest.raises(IOError):
        reproject._stream_in(p, len(data), eng.device)        # the file shrank under the reader
    assert reproject._stream_in(p, 0, eng.device).numel() == 0


def test_after_write_marker_follows_the_file(eng, tmp_path):
    """A DONE marker requested for a BED that is still being written appears only once the BED is complete on disk."""
    from salsa_b200 import reproject
    t = torch.from_numpy(np.random.default_rng(1).integers(0, 256, 3_000_000, dtype=np.uint8)).to(eng.device)
    dst, marker = str(tmp_path / 'alignment_iteration_9.bed'), str(tmp_path / 'x.DONE')
    reproject.write_behind(dst, t)
    reproject.after_write(dst, marker)
    reproject.flush(dst)
    assert os.path.exists(marker) and open(dst, 'rb').read() == bytes(t.cpu().numpy())
    assert os.path.realpath(dst) not in reproject._WRITERS      # the finished writer let go of its slot (and tensor)
    os.remove(marker)
    reproject.after_write(dst, marker)                          # nothing in flight: right away
    assert os.path.exists(marker)


def test_write_behind(eng, tmp_path):
    """background=True: the pass returns while the file is still being written; readers inside the process get the
    device copy, flush() completes the file, and a second write of the same path replaces a running one."""
    from salsa_b200 import reproject
    from oracle import py_twin
    rng = np.random.default_rng(9)
    names = ['ctg%d' % i for i in range(50)]
    lengths = [int(x) for x in rng.integers(1000, 100000, 50)]
    c2s, info = placement(rng, names, lengths, 6)
    bed = random_bed(rng, names, lengths, 40000)
    src, dst = str(tmp_path / 'alignment_iteration_1.bed'), str(tmp_path / 'alignment_iteration_2.bed')
    open(src, 'wb').write(bed)
    clen = dict(zip(names, lengths))
    run = reproject.bed_pass(eng, background=True)
    run(src, dst, c2s, info, clen, False)
    t = reproject.resident_bed(dst, eng.device, load=False)
    assert t is not None
    layout_bytes = bytes(t.cpu().numpy())
    run(src, dst, c2s, info, clen, True)                       # same path again: replaces the write in flight
    t2 = reproject.resident_bed(dst, eng.device, load=False)
    reproject.flush()
    assert not os.path.exists(dst + '.part')
    py_twin.reproject_bed(src, str(tmp_path / 'want_l.bed'), c2s, info, clen, False)
    py_twin.reproject_bed(src, str(tmp_path / 'want_r.bed'), c2s, info, clen, True)
    assert layout_bytes == open(str(tmp_path / 'want_l.bed'), 'rb').read()
    assert open(dst, 'rb').read() == open(str(tmp_path / 'want_r.bed'), 'rb').read() == bytes(t2.cpu().numpy())
    assert reproject.resident_bed(dst, eng.device, load=False) is t2          # still served from HBM after the flush


@pytest.mark.parametrize('refactor', [False, True])
@pytest.mark.parametrize('chunk', [700, 4096, 50000])
def test_chunked_pass(eng, tmp_path, refactor, chunk):
    """A BED that does not fit the GPU is re-projected in line-aligned chunks; each chunk starts with the line
    the sliding window still needs (last line / last run start of the one before).  Forced here with tiny chunks,
    over long same-contig runs that span several of them."""
    from oracle import py_twin
    from salsa_b200 import reproject
    rng = np.random.default_rng(chunk)
    names = ['ctg%d' % i for i in range(12)]
    lengths = [int(x) for x in rng.integers(1000, 100000, 12)]
    c2s, info = placement(rng, names, lengths, 4, unplaced=0.1)
    parts = [random_bed(rng, names, lengths, 300, same=0.3)]
    for k in range(6):
        parts.append(('\n'.join('ctg3\t%d\t%d\tpad%d_%d/1' % (j, j + 9, k, j) for j in range(120)) + '\n').encode())
        parts.append(random_bed(rng, names, lengths, 200, same=0.3))
    bed = b''.join(parts)
    src = str(tmp_path / 'alignment_iteration_1.bed')
    open(src, 'wb').write(bed)
    clen = dict(zip(names, lengths))
    py_twin.reproject_bed(src, str(tmp_path / 'want.bed'), c2s, info, clen, refactor)
    reproject.bed_pass(eng, chunk_bytes=chunk)(src, str(tmp_path / 'got.bed'), c2s, info, clen, refactor)
    want = open(str(tmp_path / 'want.bed'), 'rb').read()
    assert open(str(tmp_path / 'got.bed'), 'rb').read() == want and len(want) > 1000


def test_scaffold_re_proration_matches_the_host_loop(eng):
    """salsa_scaffold_re against the Python statement of layout_unitigs.py:736-814 (salsa_b200.host.reproject): random
    scaffolds incl. single contigs in both orientations, contig boundaries exactly on the midpoint, zero RE counts."""
    from salsa_b200 import reproject as gpu_reproject
    from salsa_b200.host import reproject as host
    rng = np.random.default_rng(11)
    ncontig = 4000
    names = ['ctg%05d' % i for i in range(ncontig)]
    clen = {n: int(rng.integers(1, 3_000_000)) for n in names}
    for k in range(0, 200, 2):                       # pairs of equal length: two-contig scaffolds split exactly at a boundary
        clen[names[k + 1]] = clen[names[k]]
    re_counts = {n: (int(rng.integers(0, 9000)), int(rng.integers(0, 9000))) for n in names}
    for n in names[300:330]:
        re_counts[n] = (0, 0)
    scaffolds, i, sid = {}, 0, 0
    while i < ncontig:
        k = 2 if i < 200 else int(rng.choice([1, 1, 2, 3, 5, 8, 40]))
        path = []
        for n in names[i:i + k]:
            a, b = ('B', 'E') if rng.random() < 0.5 else ('E', 'B')
            path += [n + ':' + a, n + ':' + b]
        scaffolds['scaffold_%d' % sid] = path
        i += k
        sid += 1
    c2s, info, slen = host.placement(scaffolds, clen, len)
    want = host.scaffold_re_counts(scaffolds, info, slen, clen, re_counts)
    got = gpu_reproject.re_pass(eng)(scaffolds, clen, re_counts)
    assert got == want
    assert list(got.keys()) == list(want.keys())            # file order = dict order


def test_clean_bed_pass_matches_the_host_loop(eng, tmp_path):
    """salsa_clean_bed_plan / _emit (correct.py:62-92 on the device) against the line loop of salsa_b200.host.clean:
    cut and uncut contigs, lines left of / right of / across the cut, every str.split() separator, extra columns,
    the lost tail at several buffer sizes, and the errors the reference would die with."""
    from salsa_b200 import reproject as gpu_reproject
    from salsa_b200.engine import DataError
    from salsa_b200.host import clean
    rng = np.random.default_rng(3)
    names = ['c%d' % i for i in range(40)] + ['a_rather_long_contig_name_%d' % i for i in range(5)]
    seqs = {n: 'A' * 10 for n in names}
    cut = {n: int(rng.integers(100, 5000)) for n in names[::3]}
    seps = ['\t', ' ', '  ', '\t ', '\x0b', '\r']
    lines = []
    for i in range(5000):
        c = names[int(rng.integers(len(names)))]
        s0 = int(rng.integers(0, 6000))
        e0 = s0 + int(rng.integers(1, 400))
        if c in cut and rng.random() < 0.2:
            s0, e0 = cut[c] - int(rng.integers(0, 3)), cut[c] + int(rng.integers(0, 3))      # around the cut, equality included
        sep = lambda: seps[int(rng.integers(len(seps)))]                                     # noqa: E731
        tail = ['', '\t60\t+', ' x y z'][int(rng.integers(3))]
        lines.append(c + sep() + ('+' if rng.random() < 0.02 else '') + str(s0) + sep() + str(e0) + sep() + 'r%d/%d' % (i // 2, 1 + i % 2) + tail)
    text = '\n'.join(lines) + '\n'
    src = str(tmp_path / 'in.bed')
    open(src, 'w', newline='\n').write(text)
    for flush in (1000000, 1000, 7, 1):
        want, got = str(tmp_path / 'want.bed'), str(tmp_path / 'got.bed')
        clean._clean_bed_lines(src, want, seqs, cut, flush)
        gpu_reproject.clean_pass(eng)(src, got, names, cut, flush)
        assert open(got, 'rb').read() == open(want, 'rb').read(), flush
    assert os.path.getsize(str(tmp_path / 'got.bed')) > 1000                                  # flush = 1 keeps everything
    # no trailing newline, and the errors
    for bad, code in (('c1 5 9 r/1\nc1 5 9', 1), ('c1 5 9 r/1\n\nc1 5 9 r/2\n', 1), ('zz 5 9 r/1\n', 3), ('c0 5x 9 r/1\n', 2),
                      ('c1 5x 9 r/1\n', None)):
        open(src, 'w', newline='\n').write(bad)
        if code is None:                          # c1 is not cut: its coordinates are copied as text, never parsed
            assert 'c1' not in cut
            gpu_reproject.clean_pass(eng)(src, str(tmp_path / 'g2.bed'), names, cut, 1)
            assert open(str(tmp_path / 'g2.bed')).read() == 'c1\t5x\t9\tr/1\n'
            continue
        with pytest.raises(DataError) as ge:
            gpu_reproject.clean_pass(eng)(src, str(tmp_path / 'g2.bed'), names, cut, 1)
        assert ge.value.code == code
        with pytest.raises((IndexError, KeyError, ValueError)):
            clean._clean_bed_lines(src, str(tmp_path / 'w2.bed'), seqs, cut, 1)
