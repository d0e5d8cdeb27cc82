"""Multi-GPU plumbing of the link-table path: one process per GPU, torch.distributed
(NCCL over NVLink / NVSwitch on the GPU box, gloo in the CPU tests).

The data path shards naturally -- read pairs are independent -- so the only
exchange step is the merge of the per-GPU tables:
  dense  : all-reduce(sum) of the uint64 table (288 MB for 3,000 contigs);
  sparse : every link of a rank's hash table goes into the segment of its owner
           = mix64(key) % world (one pass over the table, no compaction, no sort),
           ONE all-to-all with equal splits moves the segments, and the receiver
           adds them into its table (exchange_segments; the older compact ->
           partition -> three all-to-alls route is exchange_sparse).
Integer sums: results do not depend on the number of shards.
"""
import numpy as np
import torch
import torch.distributed as dist


def shard_ranges(view, world, involved=None):
    """Cut a BED (anything indexable as bytes: bytes, np.memmap, ...) into `world`
    byte ranges that start on line starts and overlap by exactly one line, the
    chunking rule of salsa_bed_ingest_dense (include/salsa_b200.h): shard k's first
    line is shard k-1's last line, which only plays "previous"
    (make_links.py:66-68), so every adjacent-line pair is evaluated exactly once.
    involved: with an avoid list (iterations >= 2) the set of contig names that occur
    in some avoid pair (Engine.involved, as bytes).  The reference's `continue`
    leaves prev_line stale across skipped pairs (make_links.py:72-74), so a shard
    may only start on a line that can never be skipped -- one whose contig is in no
    avoid pair; the cut walks back to such a line (as Engine.ingest_file does).
    Returns [(lo, hi)] * world; empty shards are (hi, hi)."""
    n = len(view)
    if n == 0:
        return [(0, 0)] * world
    if involved is not None:
        involved = set(x if isinstance(x, bytes) else str(x).encode('latin-1') for x in involved)

    def free_line_at_or_before(p, floor):
        # the last line start <= p whose contig is in no avoid pair (floor if there is none after it)
        while p > floor:
            tok = bytes(view[p:min(n, p + 1024)]).split(None, 1)
            if not tok or tok[0] not in involved:
                return p
            p = line_start_at_or_before(p - 1)
        return floor

    def line_start_at_or_before(p):
        # start of the line that contains byte p
        lo = max(0, p - (1 << 20))
        blk = np.frombuffer(bytes(view[lo:p]), dtype=np.uint8) if p > lo else np.zeros(0, np.uint8)
        nl = np.flatnonzero(blk == 10)
        if len(nl):
            return lo + int(nl[-1]) + 1
        if lo == 0:
            return 0
        raise ValueError('BED line longer than 1 MiB near byte %d' % p)

    cuts = [0]
    for k in range(1, world):
        p = n * k // world
        c = max(cuts[-1], line_start_at_or_before(min(p, n - 1)))
        if involved:
            c = free_line_at_or_before(c, cuts[-1])
        cuts.append(c)
    ranges = []
    for k in range(world):
        lo = cuts[k]
        if k + 1 < world:
            nxt = cuts[k + 1]
            # shard k ends after the line that STARTS at cuts[k+1] (that line is shared)
            blk = np.frombuffer(bytes(view[nxt:min(n, nxt + (1 << 20))]), dtype=np.uint8)
            nl = np.flatnonzero(blk == 10)
            hi = nxt + int(nl[0]) + 1 if len(nl) else n
        else:
            hi = n
        ranges.append((lo, hi) if hi > lo else (hi, hi))
    return ranges


def merge_dense(table):
    """Sum the per-GPU dense tables in place on every rank."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(table)
    return table


class TorchOps:
    """Reference implementations of the two device steps of the sparse merge with torch ops.
    Used by the CPU (gloo) tests of the exchange plumbing; the GPU path uses EngineOps."""

    @staticmethod
    def mix64(x):
        # must agree with salsa_mix64 (common.cuh); int64 arithmetic wraps like uint64
        def shr(v, k):
            return (v >> k) & ((1 << (64 - k)) - 1)
        x = x ^ shr(x, 33)
        x = x * torch.tensor(-49064778989728563, dtype=torch.int64)        # 0xff51afd7ed558ccd
        x = x ^ shr(x, 33)
        x = x * torch.tensor(-4265267296055464877, dtype=torch.int64)      # 0xc4ceb9fe1a85ec53
        x = x ^ shr(x, 33)
        return x

    def partition(self, keys, counts, nparts):
        h = self.mix64(keys)
        # unsigned modulo of a wrapped int64
        owner = torch.from_numpy((h.numpy().view(np.uint64) % np.uint64(nparts)).astype(np.int64))
        order = torch.argsort(owner, stable=True)
        pc = torch.bincount(owner, minlength=nparts)
        return keys[order], counts[order], pc.tolist()

    def reduce(self, keys, counts):
        if keys.numel() == 0:
            return keys, counts
        uk, inv = torch.unique(keys, return_inverse=True)
        out = torch.zeros(uk.numel(), dtype=counts.dtype)
        out.index_add_(0, inv, counts)
        return uk, out


def exchange_sparse(keys, counts, ops, group=None):
    """Hash-partitioned all-to-all of compacted (key, count) pairs followed by a local
    reduce.  keys/counts: 1-D int64 tensors (device of the process group's backend).
    Returns the (keys, counts) this rank OWNS, reduced."""
    world = dist.get_world_size(group)
    pk, pc, send = ops.partition(keys, counts, world)
    send_t = torch.tensor(send, dtype=torch.int64, device=keys.device)
    recv_t = torch.empty(world, dtype=torch.int64, device=keys.device)
    dist.all_to_all_single(recv_t, send_t, group=group)
    recv = [int(x) for x in recv_t.tolist()]
    rk = torch.empty(sum(recv), dtype=torch.int64, device=keys.device)
    rc = torch.empty(sum(recv), dtype=torch.int64, device=keys.device)
    dist.all_to_all_single(rk, pk, output_split_sizes=recv, input_split_sizes=[int(x) for x in send], group=group)
    dist.all_to_all_single(rc, pc, output_split_sizes=recv, input_split_sizes=[int(x) for x in send], group=group)
    return ops.reduce(rk, rc)


class EngineOps:
    """The device steps of the sparse merge through the C ABI (salsa_links_partition,
    salsa_sparse_add + salsa_links_compact)."""

    def __init__(self, eng):
        self.eng = eng

    def partition(self, keys, counts, nparts):
        import ctypes as C
        from ._lib import check
        L, dev = self.eng.L, self.eng.device
        n = keys.numel()
        ok = torch.empty(max(n, 1), dtype=torch.int64, device=dev)
        oc = torch.empty(max(n, 1), dtype=torch.int64, device=dev)
        scratch = torch.zeros(64, dtype=torch.int64, device=dev)
        pc = (C.c_uint64 * nparts)()
        check(L.salsa_links_partition(keys.data_ptr(), counts.data_ptr(), n, nparts, ok.data_ptr(), oc.data_ptr(),
                                      scratch.data_ptr(), pc, torch.cuda.current_stream().cuda_stream))
        return ok[:n], oc[:n], [int(x) for x in pc]

    def reduce(self, keys, counts):
        import ctypes as C
        from ._lib import check
        L, dev = self.eng.L, self.eng.device
        st = torch.cuda.current_stream().cuda_stream
        n = keys.numel()
        cap = 1 << 10
        while cap < 2 * n + 2:
            cap <<= 1
        tk = torch.empty(cap, dtype=torch.int64, device=dev)
        tc = torch.empty(cap, dtype=torch.int64, device=dev)
        check(L.salsa_sparse_reset(tk.data_ptr(), tc.data_ptr(), cap, st))
        full = torch.zeros(1, dtype=torch.int32, device=dev)
        check(L.salsa_sparse_add(keys.data_ptr(), counts.data_ptr(), n, tk.data_ptr(), tc.data_ptr(), cap, full.data_ptr(), st))
        assert int(full.item()) == 0
        scratch = torch.empty(L.salsa_links_compact_scratch_bytes(cap) // 8 + 1, dtype=torch.int64, device=dev)
        m = C.c_uint64(0)
        check(L.salsa_links_compact(tk.data_ptr(), tc.data_ptr(), cap, None, None, 0, scratch.data_ptr(), C.byref(m), st))
        ok = torch.empty(max(m.value, 1), dtype=torch.int64, device=dev)
        oc = torch.empty(max(m.value, 1), dtype=torch.int64, device=dev)
        check(L.salsa_links_compact(tk.data_ptr(), tc.data_ptr(), cap, ok.data_ptr(), oc.data_ptr(), m.value, scratch.data_ptr(),
                                    C.byref(m), st))
        return ok[:m.value], oc[:m.value]


# ---------------------------------------------------------------- segments: one pass, one all-to-all
def segment_bytes(slot_cap):
    """[u64 count][u64 pad][u64 keys[slot_cap]][u32 counts[slot_cap]]  (include/salsa_b200.h: salsa_sparse_partition)"""
    assert slot_cap % 4 == 0
    return 16 + 12 * slot_cap


class TorchSegmentOps:
    """numpy restatement of salsa_sparse_partition / salsa_sparse_add_segments for the CPU (gloo) tests of the
    exchange plumbing; the GPU path uses EngineSegmentOps."""

    def partition(self, tkeys, tcounts, nparts, slot_cap):
        k, c = tkeys.numpy(), tcounts.numpy()
        livek, livec = k[(k != -1) & (c != 0)], c[(k != -1) & (c != 0)]
        owner = (TorchOps.mix64(torch.from_numpy(livek)).numpy().view(np.uint64) % np.uint64(nparts)).astype(np.int64)
        segb = segment_bytes(slot_cap)
        buf = np.zeros(nparts * segb, np.uint8)
        overflow = False
        for p in range(nparts):
            kk, cc = livek[owner == p], livec[owner == p]
            if len(kk) > slot_cap or (len(cc) and cc.max() > 0xFFFFFFFF):
                overflow = True
                kk, cc = kk[:slot_cap], np.minimum(cc[:slot_cap], 0xFFFFFFFF)
            seg = buf[p * segb:(p + 1) * segb]
            seg[:8].view(np.uint64)[0] = len(kk)
            seg[16:16 + 8 * len(kk)].view(np.int64)[:] = kk
            seg[16 + 8 * slot_cap:16 + 8 * slot_cap + 4 * len(cc)].view(np.uint32)[:] = cc.astype(np.uint32)
        return torch.from_numpy(buf), overflow

    def add(self, recv, nparts, slot_cap):
        buf, segb = recv.numpy(), segment_bytes(slot_cap)
        acc = {}
        for p in range(nparts):
            seg = buf[p * segb:(p + 1) * segb]
            n = int(seg[:8].view(np.uint64)[0])
            kk = seg[16:16 + 8 * n].view(np.int64)
            cc = seg[16 + 8 * slot_cap:16 + 8 * slot_cap + 4 * n].view(np.uint32)
            for a, b in zip(kk.tolist(), cc.tolist()):
                acc[a] = acc.get(a, 0) + b
        ks = sorted(acc)
        return torch.tensor(ks, dtype=torch.int64), torch.tensor([acc[a] for a in ks], dtype=torch.int64)


class EngineSegmentOps:
    """salsa_sparse_partition / salsa_sparse_add_segments through the C ABI.  Buffers are kept between calls (a step of
    the merge allocates nothing after the first one); nothing here synchronises: overflow / table-full flags are device
    words that the caller reads when it wants to (flags())."""

    def __init__(self, eng, out_capacity):
        self.eng, self.outcap = eng, out_capacity
        dev = eng.device
        self.tk = torch.empty(out_capacity, dtype=torch.int64, device=dev)
        self.tc = torch.empty(out_capacity, dtype=torch.int64, device=dev)
        self.flag = torch.zeros(2, dtype=torch.int32, device=dev)        # [0] a segment overflowed, [1] owner's table full
        self.send = self.recv = None

    def _bufs(self, nbytes):
        if self.send is None or self.send.numel() != nbytes:
            self.send = torch.empty(nbytes, dtype=torch.uint8, device=self.eng.device)
            self.recv = torch.empty(nbytes, dtype=torch.uint8, device=self.eng.device)
        return self.send, self.recv

    def partition(self, tkeys, tcounts, nparts, slot_cap):
        from ._lib import check
        L = self.eng.L
        send, _ = self._bufs(nparts * L.salsa_sparse_segment_bytes(slot_cap))
        check(L.salsa_sparse_partition(tkeys.data_ptr(), tcounts.data_ptr(), tkeys.numel(), nparts, slot_cap, send.data_ptr(),
                                       self.flag[0:1].data_ptr(), torch.cuda.current_stream().cuda_stream))
        return send, None

    def add(self, recv, nparts, slot_cap):
        from ._lib import check
        L, st = self.eng.L, torch.cuda.current_stream().cuda_stream
        check(L.salsa_sparse_reset(self.tk.data_ptr(), self.tc.data_ptr(), self.outcap, st))
        check(L.salsa_sparse_add_segments(recv.data_ptr(), nparts, slot_cap, self.tk.data_ptr(), self.tc.data_ptr(), self.outcap,
                                          self.flag[1:2].data_ptr(), st))
        return self.tk, self.tc

    def flags(self):
        """(segment overflow, owner table full) -- synchronises."""
        f = self.flag.cpu().tolist()
        return bool(f[0]), bool(f[1])


def exchange_segments(tkeys, tcounts, ops, slot_cap, group=None):
    """The sparse merge in one pass and one collective: `tkeys`/`tcounts` are a rank's whole open-addressing table
    (empty slots: key -1); every live link is appended to the segment of owner = mix64(key) % world, the segments are
    exchanged by ONE all-to-all with equal splits (no sizes to agree on, so no host round trip), and the receiver adds
    what it got into its table.  Returns what ops.add returns: the links this rank owns."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    send, _ = ops.partition(tkeys, tcounts, world, slot_cap)
    if world > 1:
        recv = ops.recv if getattr(ops, 'recv', None) is not None else torch.empty_like(send)
        dist.all_to_all_single(recv, send, group=group)
    else:
        recv = send
    return ops.add(recv, world, slot_cap)


# ---------------------------------------------------------------- RE-site counting, sharded by contig
def fasta_shard_ranges(view, world):
    """Cut a FASTA (anything indexable as bytes) into `world` byte ranges that each start on a record header ('>' at a
    line start) and are balanced by bytes, i.e. by bases: records are independent units (RE_sites.py:30-88 counts per
    contig), so a rank runs the whole single-GPU path on its range.  Returns [(lo, hi)] * world; empty shards are (hi, hi)."""
    n = len(view)
    if n == 0:
        return [(0, 0)] * world

    def header_at_or_after(p):
        if p <= 0:
            return 0
        step = 1 << 22
        q = p - 1                                         # a header starts right after a newline
        while q < n:
            blk = np.frombuffer(bytes(view[q:min(n, q + step)]), dtype=np.uint8)
            hit = np.flatnonzero((blk[:-1] == 10) & (blk[1:] == 62))
            if len(hit):
                return q + int(hit[0]) + 1
            q += step - 1                                 # overlap one byte: a "\n>" may straddle the blocks
        return n

    cuts = [0]
    for k in range(1, world):
        cuts.append(max(cuts[-1], header_at_or_after(n * k // world)))
    cuts.append(n)
    return [(cuts[k], cuts[k + 1]) if cuts[k + 1] > cuts[k] else (cuts[k + 1], cuts[k + 1]) for k in range(world)]


def merge_site_counts(parts):
    """parts: per rank (names, left, right, seqlen) in rank order = file order.  The reference fills dicts keyed by
    record name (RE_sites.py:19,61): a repeated name keeps its first position and its last sequence."""
    order, last = [], {}
    for names, left, right, seqlen in parts:
        for i, nm in enumerate(names):
            if nm not in last:
                order.append(nm)
            last[nm] = (int(left[i]), int(right[i]), int(seqlen[i]))
    left = np.array([last[nm][0] for nm in order], dtype=np.int64)
    right = np.array([last[nm][1] for nm in order], dtype=np.int64)
    seqlen = np.array([last[nm][2] for nm in order], dtype=np.int64)
    return order, left, right, seqlen


def count_sites_sharded(count_fn, fasta, enzyme, group=None):
    """RE-site counting over all ranks of the process group: rank r counts the records of its byte range with
    count_fn(shard, enzyme) -> (names, left, right, seqlen) (salsa_b200.re_sites.count_sites bound to an Engine; the CPU
    tests pass the oracle's), the per-contig results -- two integers per contig -- are all-gathered, and every rank
    returns the whole table.  An error on any rank (a ':' in a header, a bad enzyme) is raised on every rank."""
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    lo, hi = fasta_shard_ranges(fasta, world)[rank]
    mine, err = ([], [], [], []), None
    try:
        if hi > lo:
            names, left, right, seqlen = count_fn(fasta[lo:hi], enzyme)
            mine = (list(names), [int(x) for x in left], [int(x) for x in right], [int(x) for x in seqlen])
        elif rank == 0:
            mine = tuple(list(x) for x in count_fn(fasta[0:0], enzyme))
    except Exception as e:           # noqa: BLE001  (re-raised below, on every rank)
        err = e
    if world > 1:
        parts = [None] * world
        dist.all_gather_object(parts, (mine, None if err is None else (type(err).__name__, str(err))), group=group)
    else:
        parts = [(mine, None if err is None else (type(err).__name__, str(err)))]
    for r, (_, e) in enumerate(parts):
        if e is not None:
            if err is not None:
                raise err
            raise RuntimeError('rank %d: %s: %s' % (r, e[0], e[1]))
    return merge_site_counts([p for p, _ in parts])


# ---------------------------------------------------------------- coverage (misassembly detection), sharded by contig
def shard_lengths(lengths_text, world, min_len=1_000_000):
    """Deal the lines of a lengths file (`name\tlen...`) to `world` ranks: contigs of at least `min_len` bases -- the only
    ones the coverage programs look at (break_contigs_start.cpp:325, break_contigs.cpp:218-225) -- longest first to the
    least loaded rank; shorter ones all to rank 0 (they cost nothing).  Every rank's text keeps the file's line order.
    Returns a list of `world` texts (bytes)."""
    lines = [ln for ln in lengths_text.split(b'\n') if ln.strip()]
    big = []
    for i, ln in enumerate(lines):
        a = ln.split()
        try:
            v = int(float(a[1]))
        except (IndexError, ValueError):
            v = 0
        big.append((v, i))
    load, owner = [0] * world, [0] * len(lines)
    for v, i in sorted(big, key=lambda t: (-t[0], t[1])):
        if v >= min_len:
            r = min(range(world), key=lambda k: (load[k], k))
            owner[i] = r
            load[r] += v
    return [b''.join(lines[i] + b'\n' for i in range(len(lines)) if owner[i] == r) for r in range(world)]


def merge_coverage_reports(parts, joins):
    """Per-rank stdout of break_contigs_start (joins = False: `contig\tposition` lines) or break_contigs (joins = True:
    `scaffold(\tpos)*` lines + the three summary lines of break_contigs.cpp:353-355) -> one report.  The per-contig lines
    are independent; the summary counts add up and the percentage is recomputed as the reference prints it
    (cout << double: 6 significant digits; 0/0 prints -nan)."""
    body, total, susp = [], 0, 0
    for text in parts:
        for ln in text.split(b'\n'):
            if not ln:
                continue
            if joins and ln.startswith(b'Total Joins = '):
                total += int(ln.split(b'=')[1])
            elif joins and ln.startswith(b'Suspicious Joins = '):
                susp += int(ln.split(b'=')[1])
            elif joins and ln.startswith(b'Percent Suspicious Joins = '):
                pass
            else:
                body.append(ln)
    out = b''.join(ln + b'\n' for ln in body)
    if joins:
        pct = b'-nan' if total == 0 else ('%g' % (susp * 100.0 / total)).encode()
        out += b'Total Joins = %d\nSuspicious Joins = %d\nPercent Suspicious Joins = %s\n' % (total, susp, pct)
    return out


def coverage_sharded(run_fn, lengths_path, joins, workdir, group=None):
    """The coverage programs over all ranks: contigs are independent units (one difference array each), so rank r runs
    the whole program -- run_fn(lengths_path_r) -> stdout bytes -- on ITS contigs only (its lines of the lengths file;
    every other contig then counts as unknown, which the programs skip exactly like a short one) and the reports are
    gathered and merged.  Every rank scans the whole BED (it is resident on every GPU anyway; the scan is 12 of 90 ms at
    BASELINE size), the prefix/Kadane analysis -- the bulk -- is split.  Returns the merged report on every rank."""
    import os
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    with open(lengths_path, 'rb') as f:
        texts = shard_lengths(f.read(), world)
    mine = os.path.join(workdir, 'lengths.rank%d' % rank)
    with open(mine, 'wb') as f:
        f.write(texts[rank])
    text = run_fn(mine)
    if world > 1:
        parts = [None] * world
        dist.all_gather_object(parts, text, group=group)
    else:
        parts = [text]
    return merge_coverage_reports(parts, joins)
