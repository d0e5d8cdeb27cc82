"""Synthetic assemblies and Hi-C BEDs of the shapes BASELINE.json names.

Written with torch tensor ops only, so the same code generates small cases on
the CPU (tests, fixtures) and multi-GB shards directly in HBM (bench.py; a
45 GB BED does not fit this container's disk budget or PCIe patience).  The
generator knows every record's contig id and coordinates, so it can also state
the EXPECTED link table without ever parsing text (``expected_links``) -- the
full-size parity property used by bench.py.

Recipe: contigs with distinct log-uniform lengths are
laid along hidden chromosomes; a read pair is a genome position x and a mate
at distance d (P(d) ~ 1/d, i.e. log-uniform) or, for `noise` of the pairs, a
uniformly random position; both ends are mapped back to contig coordinates.
`triple` of the reads get a third record (a second alignment of mate 1 on a
random contig) to exercise make_links.py's sliding adjacent-line window
(make_links.py:60-146).  BED text is the 6-column bamToBed form
``ctgNNNNNN\\tstart\\tend\\treadNNNNNNNNNN/1\\t60\\t+``; records are in read-name
order because read ids are zero padded and sequential (README.md:68-76).
"""
import math

import numpy as np
import torch

READ_LEN = 150

CONFIGS = {
    # BASELINE.json configs[0..4]; lengths bounds follow SURVEY.md 8(d)
    'cfg1': dict(ncontig=200, genome=2_000_000, len_lo=2_000, len_hi=40_000, pairs=1_000_000, nchrom=4),
    'cfg2': dict(ncontig=3_000, genome=3_100_000_000, len_lo=50_000, len_hi=20_000_000, pairs=500_000_000, nchrom=23),
    'cfg3': dict(ncontig=3_000, genome=3_100_000_000, len_lo=50_000, len_hi=20_000_000, pairs=500_000_000, nchrom=23),
    'cfg4': dict(ncontig=3_000, genome=3_100_000_000, len_lo=50_000, len_hi=20_000_000, pairs=1_000_000_000, nchrom=23),
    'cfg5': dict(ncontig=100_000, genome=30_000_000_000, len_lo=20_000, len_hi=5_000_000, pairs=4_000_000_000, nchrom=12),
}


class Assembly:
    """names[i] = b'ctg%06d' % i (so table order == byte-sorted order)."""

    def __init__(self, lengths, order, orient, nchrom, seed):
        self.lengths = np.asarray(lengths, dtype=np.int64)
        self.ncontig = len(self.lengths)
        self.names = [b'ctg%06d' % i for i in range(self.ncontig)]
        self.order = np.asarray(order, dtype=np.int64)      # hidden genome order -> contig id
        self.orient = np.asarray(orient, dtype=np.int64)    # per contig id: 0 forward, 1 reversed
        self.nchrom = nchrom
        self.seed = seed
        # hidden genome coordinate of each contig's first base (in hidden order)
        olen = self.lengths[self.order]
        self.gstart_ordered = np.concatenate([[0], np.cumsum(olen)])          # ncontig+1
        self.genome = int(self.gstart_ordered[-1])
        # chromosome boundaries: contiguous groups of the hidden order
        cuts = np.linspace(0, self.ncontig, nchrom + 1).astype(np.int64)
        self.chrom_of_ordered = np.zeros(self.ncontig, dtype=np.int64)
        for k in range(nchrom):
            self.chrom_of_ordered[cuts[k]:cuts[k + 1]] = k

    def lengths_text(self):
        """scaffold_length_iteration_1 / .fai-like: name, len (+ ignored columns)."""
        return b''.join(b'%s\t%d\t0\t60\t61\n' % (n, l) for n, l in zip(self.names, self.lengths))


def make_assembly(ncontig, genome, len_lo, len_hi, nchrom=4, seed=20261106, **_):
    rng = np.random.Generator(np.random.Philox(seed))
    ln = np.exp(rng.uniform(math.log(len_lo), math.log(len_hi), ncontig))
    ln = ln * (genome / ln.sum())
    ln = np.maximum(np.round(ln).astype(np.int64), max(READ_LEN * 2, 400))
    # distinct lengths (get_seq.py:114 sorts by length; ties would be order dependent)
    ln = np.sort(ln)
    for i in range(1, ncontig):
        if ln[i] <= ln[i - 1]:
            ln[i] = ln[i - 1] + 1
    rng.shuffle(ln)
    order = rng.permutation(ncontig)
    orient = rng.integers(0, 2, ncontig)
    return Assembly(ln, order, orient, nchrom, seed)


def gen_fasta(asm, seed=None, n_rate=0.001, width=60, enzyme_hint=None):
    """Host FASTA text for small assemblies: uniform ACGT, a few N runs, one
    lower-case block per 100 contigs (sequence N / lower case never match,
    RE_sites.py:66)."""
    rng = np.random.Generator(np.random.Philox((seed if seed is not None else asm.seed) + 7))
    out = []
    alpha = np.frombuffer(b'ACGT', dtype=np.uint8)
    for i, (nm, L) in enumerate(zip(asm.names, asm.lengths)):
        s = alpha[rng.integers(0, 4, int(L))]
        nrun = rng.poisson(L * n_rate / 20.0)
        for _ in range(nrun):
            p = int(rng.integers(0, L))
            s[p:p + int(rng.integers(1, 40))] = ord('N')
        if i % 100 == 0 and L > 200:
            p = int(rng.integers(0, L - 100))
            s[p:p + 100] = np.frombuffer(bytes(s[p:p + 100]).lower(), dtype=np.uint8)
        out.append(b'>' + nm + b' len=%d\n' % L)
        body = bytes(s)
        out.append(b'\n'.join(body[j:j + width] for j in range(0, len(body), width)) + b'\n')
    return b''.join(out)


def _t(a, device):
    return torch.as_tensor(np.asarray(a), device=device)


def gen_records(asm, npairs, seed, device='cpu', read_id_base=0, noise=0.05, triple=0.005,
                dmin=300.0, dmax=None):
    """Returns per-LINE tensors (int64): ctg, start, end, read, mate (1|2), in
    read-name order.  Pair i contributes lines (r/1, [r/1 extra], r/2)."""
    g = torch.Generator(device=device)
    g.manual_seed(int(seed))
    dev = torch.device(device)
    G = asm.genome
    if dmax is None:
        dmax = float(max(4 * np.median(asm.lengths), 2000.0))
    u = torch.rand(npairs, generator=g, device=dev, dtype=torch.float64)
    x = (u * (G - 1)).to(torch.int64)
    ud = torch.rand(npairs, generator=g, device=dev, dtype=torch.float64)
    d = torch.exp(math.log(dmin) + ud * (math.log(dmax) - math.log(dmin))).to(torch.int64)
    sign = torch.randint(0, 2, (npairs,), generator=g, device=dev) * 2 - 1
    y = x + sign * d
    un = torch.rand(npairs, generator=g, device=dev)
    yn = (torch.rand(npairs, generator=g, device=dev, dtype=torch.float64) * (G - 1)).to(torch.int64)
    y = torch.where(un < noise, yn, y)
    y = y.clamp_(0, G - 1)
    gstart = _t(asm.gstart_ordered, dev)
    order = _t(asm.order, dev)
    orient = _t(asm.orient, dev)
    lengths = _t(asm.lengths, dev)
    chrom = _t(asm.chrom_of_ordered, dev)

    def locate(p):
        k = torch.searchsorted(gstart, p, right=True) - 1          # hidden slot
        k = k.clamp_(0, asm.ncontig - 1)
        c = order[k]
        off = p - gstart[k]
        L = lengths[c]
        loc = torch.where(orient[c] == 1, L - 1 - off, off)
        s = (loc - READ_LEN // 2).clamp_(min=0)
        s = torch.minimum(s, L - READ_LEN)
        return c, s, s + READ_LEN, k

    c1, s1, e1, k1 = locate(x)
    c2, s2, e2, k2 = locate(y)
    # pairs crossing a chromosome boundary become noise pairs (kept: still valid records)
    del k1, k2, chrom
    # third record for `triple` of the reads: mate 1 aligned a second time elsewhere
    ut = torch.rand(npairs, generator=g, device=dev)
    has3 = ut < triple
    z = (torch.rand(npairs, generator=g, device=dev, dtype=torch.float64) * (G - 1)).to(torch.int64)
    c3, s3, e3, _ = locate(z)
    nl = 2 + has3.to(torch.int64)
    off = torch.cumsum(nl, 0) - nl
    n = int(nl.sum().item())
    ctg = torch.empty(n, dtype=torch.int64, device=dev)
    st = torch.empty(n, dtype=torch.int64, device=dev)
    en = torch.empty(n, dtype=torch.int64, device=dev)
    rd = torch.empty(n, dtype=torch.int64, device=dev)
    mt = torch.empty(n, dtype=torch.int64, device=dev)
    rid = torch.arange(npairs, device=dev, dtype=torch.int64) + read_id_base
    ctg[off] = c1; st[off] = s1; en[off] = e1; rd[off] = rid; mt[off] = 1
    i3 = off[has3] + 1
    ctg[i3] = c3[has3]; st[i3] = s3[has3]; en[i3] = e3[has3]; rd[i3] = rid[has3]; mt[i3] = 1
    last = off + nl - 1
    ctg[last] = c2; st[last] = s2; en[last] = e2; rd[last] = rid; mt[last] = 2
    return dict(ctg=ctg, start=st, end=en, read=rd, mate=mt)


def _ndigits(v):
    nd = torch.ones_like(v)
    p = 10
    for _ in range(18):
        nd += (v >= p).to(v.dtype)
        p *= 10
    return nd


def format_bed(rec, out=None):
    """Records -> BED text as a uint8 tensor on the records' device."""
    ctg, st, en, rd, mt = rec['ctg'], rec['start'], rec['end'], rec['read'], rec['mate']
    dev = ctg.device
    n = ctg.numel()
    ds, de = _ndigits(st), _ndigits(en)
    # ctgNNNNNN \t S \t E \t readNNNNNNNNNN / m \t 6 0 \t + \n
    ll = 9 + 1 + ds + 1 + de + 1 + 14 + 2 + 1 + 2 + 1 + 1 + 1
    off = torch.cumsum(ll, 0) - ll
    total = int(ll.sum().item()) if n else 0
    buf = out if out is not None else torch.empty(total, dtype=torch.uint8, device=dev)
    assert buf.numel() >= total

    def put_const(pos, s):
        for j, ch in enumerate(s):
            buf[pos + j] = ch

    def put_fixed(pos, val, width):
        v = val.clone()
        for j in range(width - 1, -1, -1):
            buf[pos + j] = (48 + v % 10).to(torch.uint8)
            v //= 10

    def put_var(pos, val, nd, maxd=11):
        v = val.clone()
        for k in range(maxd):
            m = nd > k
            if not bool(m.any()):
                break
            buf[(pos + nd - 1 - k)[m]] = (48 + v[m] % 10).to(torch.uint8)
            v //= 10

    p = off
    put_const(p, b'ctg'); put_fixed(p + 3, ctg, 6); put_const(p + 9, b'\t')
    p = p + 10
    put_var(p, st, ds); p = p + ds; put_const(p, b'\t'); p = p + 1
    put_var(p, en, de); p = p + de; put_const(p, b'\t'); p = p + 1
    put_const(p, b'read'); put_fixed(p + 4, rd, 10); put_const(p + 14, b'/')
    buf[p + 15] = (48 + mt).to(torch.uint8)
    put_const(p + 16, b'\t60\t')
    buf[p + 20] = torch.where((rd + mt) % 2 == 0, 43, 45).to(torch.uint8)   # '+' / '-'
    put_const(p + 21, b'\n')
    return buf[:total]


def expected_links(rec, lengths, ncontig):
    """The link table make_links.py would produce from these records (no avoid
    list), computed from ids/coordinates, never from text:
    adjacent lines, different contig, same read -> key (make_links.py:71-144).
    Returns (keys, counts) with key = ((first*ncontig + second) << 2) | ends,
    sorted by key."""
    ctg, st, en, rd = rec['ctg'], rec['start'], rec['end'], rec['read']
    dev = ctg.device
    L = torch.as_tensor(np.asarray(lengths), device=dev).to(torch.int64)
    a, b = ctg[:-1], ctg[1:]
    m = (a != b) & (rd[:-1] == rd[1:])
    a, b = a[m], b[m]
    sa = (st[:-1] + en[:-1])[m]
    sb = (st[1:] + en[1:])[m]
    first = torch.minimum(a, b)
    second = torch.maximum(a, b)
    sf = torch.where(a < b, sa, sb)
    ss = torch.where(a < b, sb, sa)
    ef = (sf > L[first]).to(torch.int64)      # B iff start+end <= len
    es = (ss > L[second]).to(torch.int64)
    key = ((first * ncontig + second) << 2) | (ef * 2 + es)
    keys, counts = torch.unique(key, return_counts=True)
    return keys, counts


def small_case(ncontig=12, genome=200_000, npairs=3000, seed=1, len_lo=2_000, len_hi=60_000, nchrom=2, **kw):
    """Convenience for tests: (assembly, records, bed bytes)."""
    asm = make_assembly(ncontig, genome, len_lo, len_hi, nchrom=nchrom, seed=seed)
    rec = gen_records(asm, npairs, seed + 1, **kw)
    bed = bytes(format_bed(rec).numpy())
    return asm, rec, bed
