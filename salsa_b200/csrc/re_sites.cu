// re_sites.cu -- restriction-site counting over bit-plane packed contigs.
//
// Replaces RE_sites.py:6-28 (parse_fasta) and :45-88 (per contig, per expanded
// enzyme pattern, non-overlapping re.finditer matches split at len/2).
//
// Pipeline (all on the device, FASTA text resident in HBM):
//   1. fasta_headers   : positions of '>' at a line start (record boundaries)
//   2. fasta_count     : per 4 KiB text tile, number of KEPT sequence bytes
//                        (not in a header line, not '\n', not trailing
//                        whitespace -- ln.rstrip(), RE_sites.py:22)
//   3. (scan of tile counts: links_table.cu's single-block scan)
//   4. fasta_pack      : kept bytes -> three bit planes per 64-base block
//                        (hi, lo, valid); A=00 C=01 G=10 T=11, valid = 0 for
//                        anything else (N, lower case, stray characters never
//                        match a literal [ACGT]+ pattern, RE_sites.py:66)
//   5. re_count        : all patterns in one pass over the planes, bit-parallel
//                        across 64 positions; matches whose start < len/2 are
//                        "left" (RE_sites.py:69-74)
// Patterns with a self-overlap border (e.g. AAAA) need re.finditer's leftmost
// non-overlapping rule; for those the occurrence bit masks are written out and
// a chain-walking kernel selects the matches.
#include "common.cuh"

namespace {

constexpr int FT = 4096;        // text tile
constexpr int FHALO = 1024;     // look-ahead for trailing whitespace
constexpr int FTHREADS = 256;

__device__ __forceinline__ bool is_ws_b(uint32_t c) { return c == 32u || (c - 9u) <= 4u; }

// ---- exact SWAR byte tests (0x80 in every byte that satisfies the test) and 16-flag packing ----
__device__ __forceinline__ uint32_t eq4(uint32_t w, uint32_t c4)
{
    uint32_t x = w ^ c4;
    uint32_t t = (x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu;
    return ~(t | x) & 0x80808080u;
}
__device__ __forceinline__ uint32_t le20_4(uint32_t w)
{
    uint32_t t = (w & 0x7F7F7F7Fu) + 0x5F5F5F5Fu;
    return ~(t | w) & 0x80808080u;
}
__device__ __forceinline__ uint32_t pack16f(uint32_t z0, uint32_t z1, uint32_t z2, uint32_t z3)
{
    uint32_t lo = __dp4a(z0, 0x08040201u, 0u);
    lo = __dp4a(z1, 0x80402010u, lo);
    uint32_t hi = __dp4a(z2, 0x08040201u, 0u);
    hi = __dp4a(z3, 0x80402010u, hi);
    return (lo >> 7) | (hi << 1);
}

// ---- 1. header positions ------------------------------------------------
// 16 bytes per thread and load; a '>' is a header only at the start of the text or right after a newline
__global__ void fasta_headers_kernel(const uint8_t *txt, size_t n, unsigned long long *out, unsigned int *count,
                                     unsigned int cap)
{
    const size_t n16 = n / 16;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const bool aligned = ((uintptr_t)txt & 15u) == 0;
    for (; i < n16; i += stride) {
        uint32_t m;
        if (aligned) {
            const uint4 v = __ldg(reinterpret_cast<const uint4 *>(txt) + i);
            m = pack16f(eq4(v.x, 0x3E3E3E3Eu), eq4(v.y, 0x3E3E3E3Eu), eq4(v.z, 0x3E3E3E3Eu), eq4(v.w, 0x3E3E3E3Eu));
        } else {
            m = 0;
            for (int j = 0; j < 16; j++) m |= (txt[i * 16 + j] == '>') << j;
        }
        while (m) {
            const int b = __ffs(m) - 1;
            m &= m - 1;
            const size_t p = i * 16 + b;
            if (p == 0 || txt[p - 1] == '\n') {
                unsigned int k = atomicAdd(count, 1u);
                if (k < cap) out[k] = (unsigned long long)p;
            }
        }
    }
    if (blockIdx.x == 0 && threadIdx.x < n - n16 * 16) {          // the last, partial 16 bytes
        const size_t p = n16 * 16 + threadIdx.x;
        if (txt[p] == '>' && (p == 0 || txt[p - 1] == '\n')) {
            unsigned int k = atomicAdd(count, 1u);
            if (k < cap) out[k] = (unsigned long long)p;
        }
    }
}

// Sequence spans: span r = [sbeg[r], send[r]) is the text between the end of
// header line r and the start of the next header (or EOF).  A byte at p in a
// span is KEPT unless it is '\n' or whitespace followed only by whitespace up
// to the next '\n' / span end.
struct Spans {
    const unsigned long long *sbeg, *send;   // sorted, disjoint
    int nspan;
};

__device__ __forceinline__ int span_of(const Spans &S, unsigned long long p)
{
    // last span with sbeg <= p
    int lo = 0, hi = S.nspan;
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (S.sbeg[mid] <= p) lo = mid + 1; else hi = mid;
    }
    return lo - 1;
}

// Spans that can intersect one tile (+ halo), cached in shared memory: a per-byte binary
// search over the global span table would cost a dozen dependent global loads per byte.
constexpr int SC = 16;
struct SpanCache {
    int r0, n;                       // first cached span index; n < 0: too many spans in this tile, search globally
    unsigned long long beg[SC], end[SC];
};

__device__ void span_cache_fill(const Spans &S, size_t t0, size_t t1, SpanCache *c)
{
    int r = span_of(S, t0);
    if (r < 0) r = 0;
    c->r0 = r;
    int k = 0;
    while (r < S.nspan && S.sbeg[r] < t1) {
        if (k == SC) { k = -1; break; }
        c->beg[k] = S.sbeg[r];
        c->end[k] = S.send[r];
        k++; r++;
    }
    c->n = k;
}

// span containing-or-preceding p: returns its index (or -1) and its end
__device__ __forceinline__ int span_find(const Spans &S, const SpanCache &c, unsigned long long p, unsigned long long *e)
{
    if (c.n < 0) {
        int r = span_of(S, p);
        *e = r >= 0 ? S.send[r] : 0;
        return r;
    }
    int hit = -1;
    for (int i = 0; i < c.n; i++)
        if (c.beg[i] <= p) hit = i;
    if (hit < 0) { *e = 0; return -1; }
    *e = c.end[hit];
    return c.r0 + hit;
}

// Stage [t0, t0 + FT + FHALO) of the text into smem and compute, for the tile's
// FT bytes, kept flags (1 byte each, in `kept`).  Returns via *err when a
// whitespace run cannot be resolved inside the halo.
__device__ void tile_kept(const uint8_t *txt, size_t n, const Spans &S, size_t t0, uint8_t *sm, uint8_t *kept, int *err,
                          SpanCache *sc)
{
    const int tid = threadIdx.x;
    const int tot = FT + FHALO;
    if (tid == 0) span_cache_fill(S, t0, t0 + tot, sc);
    for (int i = tid; i < tot; i += FTHREADS) {
        size_t p = t0 + i;
        sm[i] = p < n ? txt[p] : (uint8_t)'\n';
    }
    __syncthreads();
    // span membership per byte: a byte is sequence text iff inside some span
    for (int i = tid; i < FT; i += FTHREADS) {
        size_t p = t0 + i;
        uint8_t k = 0;
        if (p < n) {
            unsigned long long send_r;
            int r = span_find(S, *sc, p, &send_r);
            if (r >= 0 && p < send_r) {
                uint32_t c = sm[i];
                if (c != '\n') {
                    k = 1;
                    if (is_ws_b(c)) {
                        // trailing iff only whitespace until the next '\n' or the span end
                        size_t e = send_r;
                        int j = i + 1;
                        for (;;) {
                            size_t q = t0 + j;
                            if (q >= e) { k = 0; break; }
                            if (j >= tot) { atomicExch(err, 1); break; }
                            uint32_t d = sm[j];
                            if (d == '\n') { k = 0; break; }
                            if (!is_ws_b(d)) break;
                            j++;
                        }
                    }
                }
            }
        }
        kept[i] = k;
    }
    __syncthreads();
}

// PLAIN tiles -- the whole 4 KiB tile lies inside ONE span and holds no whitespace but newlines (every tile of an
// ordinary FASTA except the few with a header line) -- need none of the above: a byte is kept iff it is not '\n'.
// Each thread takes 16 bytes with one load; returns true (block-uniform) when the tile is plain, with v = the thread's
// bytes, kept16 = its kept mask, and r = the span.  Falls back (false) on anything else, misaligned text included.
__device__ bool tile_plain(const uint8_t *txt, size_t n, const Spans &S, size_t t0, uint4 &v, uint32_t &kept16, int &r)
{
    __shared__ int s_r;
    if (t0 + FT > n || ((uintptr_t)txt & 15u)) return false;                   // (block-uniform conditions)
    if (threadIdx.x == 0) {
        int q = span_of(S, t0);
        s_r = (q >= 0 && S.sbeg[q] <= t0 && S.send[q] >= t0 + FT) ? q : -1;
    }
    v = __ldg(reinterpret_cast<const uint4 *>(txt + t0) + threadIdx.x);
    const uint32_t nl = pack16f(eq4(v.x, 0x0A0A0A0Au), eq4(v.y, 0x0A0A0A0Au), eq4(v.z, 0x0A0A0A0Au), eq4(v.w, 0x0A0A0A0Au));
    const uint32_t lo = pack16f(le20_4(v.x), le20_4(v.y), le20_4(v.z), le20_4(v.w));    // every byte <= 0x20 (a superset of str whitespace)
    kept16 = ~nl & 0xFFFFu;
    const int odd = __syncthreads_or((lo & ~nl) != 0u);                         // also publishes s_r
    r = s_r;
    return r >= 0 && !odd;
}

__global__ void __launch_bounds__(FTHREADS) fasta_count_kernel(const uint8_t *txt, size_t n, Spans S, uint64_t *tile_cnt, int *err)
{
    __shared__ uint8_t sm[FT + FHALO];
    __shared__ uint8_t kept[FT];
    __shared__ int red[FTHREADS / 32];
    size_t t0 = (size_t)blockIdx.x * FT;
    __shared__ SpanCache sc;
    {
        uint4 v;
        uint32_t k16;
        int r;
        if (tile_plain(txt, n, S, t0, v, k16, r)) {
            int c = __popc(k16);
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
            if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = c;
            __syncthreads();
            if (threadIdx.x == 0) {
                int t = 0;
                for (int w = 0; w < FTHREADS / 32; w++) t += red[w];
                tile_cnt[blockIdx.x] = (uint64_t)t;
            }
            return;
        }
        __syncthreads();
    }
    tile_kept(txt, n, S, t0, sm, kept, err, &sc);
    int c = 0;
    for (int i = threadIdx.x; i < FT; i += FTHREADS) c += kept[i];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < FTHREADS / 32; w++) t += red[w];
        tile_cnt[blockIdx.x] = (uint64_t)t;
    }
}

// kept-byte prefix P(p) = tile_excl[p / FT] + (kept bytes in the tile before p).
// One CTA per span computes P(sbeg[r]) and P(send[r]) -> span length.
__global__ void __launch_bounds__(FTHREADS) span_prefix_kernel(const uint8_t *txt, size_t n, Spans S, const uint64_t *tile_excl,
                                                               uint64_t *pbeg, uint64_t *plen, int *err)
{
    __shared__ uint8_t sm[FT + FHALO];
    __shared__ uint8_t kept[FT];
    __shared__ int red[FTHREADS / 32];
    __shared__ SpanCache sc;
    int r = blockIdx.x;
    uint64_t res[2];
    for (int which = 0; which < 2; which++) {
        unsigned long long p = which ? S.send[r] : S.sbeg[r];
        size_t t = (size_t)(p / FT);
        size_t t0 = t * FT;
        int upto = (int)(p - t0);
        int c = 0;
        if (upto > 0) {
            tile_kept(txt, n, S, t0, sm, kept, err, &sc);
            for (int i = threadIdx.x; i < upto; i += FTHREADS) c += kept[i];
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
        __syncthreads();
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = c;
        __syncthreads();
        int tsum = 0;
        for (int w = 0; w < FTHREADS / 32; w++) tsum += red[w];
        res[which] = tile_excl[t] + (uint64_t)tsum;     // tile_excl has ntiles + 1 entries
        __syncthreads();
    }
    if (threadIdx.x == 0) { pbeg[r] = res[0]; plen[r] = res[1] - res[0]; }
}

// ---- 4. pack ---------------------------------------------------------------
// planes: per 64-base block three uint64: [hi, lo, valid].  Record r starts at
// block blk_off[r]; base i of record r is bit (i & 63) of block blk_off[r] + i/64.
__global__ void __launch_bounds__(FTHREADS) fasta_pack_kernel(const uint8_t *txt, size_t n, Spans S, const uint64_t *tile_excl,
                                                              const uint64_t *pbeg, const uint64_t *blk_off,
                                                              unsigned long long *planes, int *err)
{
    __shared__ uint8_t sm[FT + FHALO];
    __shared__ uint8_t kept[FT];
    __shared__ uint8_t code[FT];        // compacted: code | valid << 2 | segment-change flag via seg arrays
    __shared__ unsigned long long opos[FT];   // output base index (global, in bases) of each compacted byte
    __shared__ int wtot[FTHREADS / 32 + 1];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    size_t t0 = (size_t)blockIdx.x * FT;
    __shared__ SpanCache sc;
    {
        // PLAIN tile: the kept bytes are one contiguous piece of one record's sequence.  Every thread turns its 16 bytes
        // into 3-bit codes (valid << 2 | 2-bit base), compacts the kept ones into `code` at its scan offset, and then 32
        // bases at a time are assembled into one 32-bit half of the three plane words: a plain store when the half lies
        // wholly inside this tile's piece, an atomicOr at the piece's two ragged ends (the planes start zeroed).
        uint4 v;
        uint32_t k16;
        int r;
        if (tile_plain(txt, n, S, t0, v, k16, r)) {
            const int c = __popc(k16);
            int inc = c;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                int o = __shfl_up_sync(0xffffffffu, inc, d);
                if (lane >= d) inc += o;
            }
            if (lane == 31) wtot[warp] = inc;
            __syncthreads();
            if (tid == 0) {
                int q = 0;
                for (int w = 0; w < FTHREADS / 32; w++) { int x = wtot[w]; wtot[w] = q; q += x; }
                wtot[FTHREADS / 32] = q;
            }
            __syncthreads();
            int pos = wtot[warp] + inc - c;
            const uint32_t ws4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int j = 0; j < 16; j++) {
                if (!((k16 >> j) & 1u)) continue;
                const uint32_t ch = (ws4[j >> 2] >> (8 * (j & 3))) & 255u;
                const uint32_t x = (ch >> 1) & 3u;                                        // A 0, C 1, T 2, G 3
                const uint32_t want = (0x47544341u >> (8u * x)) & 255u;                   // 'A','C','T','G'
                code[pos++] = (uint8_t)(ch == want ? ((x ^ (x >> 1)) | 4u) : 0u);            // A 00 C 01 G 10 T 11; anything else: no bits
            }
            __syncthreads();
            const int k = wtot[FTHREADS / 32];
            const unsigned long long K = blk_off[r] * 64ull + (tile_excl[blockIdx.x] - pbeg[r]);   // output base index of code[0]
            uint32_t *planes32 = reinterpret_cast<uint32_t *>(planes);
            const unsigned long long h0 = K >> 5;                                          // first 32-base half touched
            const int nh = (int)(((K + (unsigned long long)k + 31ull) >> 5) - h0);
            for (int hidx = tid; hidx < nh; hidx += FTHREADS) {
                const unsigned long long hb = h0 + (unsigned long long)hidx;               // half index: block hb >> 1, half hb & 1
                const long long p0 = (long long)(hb << 5) - (long long)K;                  // code index of this half's bit 0
                uint32_t hi = 0, lo = 0, va = 0;
                const int a = p0 < 0 ? (int)-p0 : 0, b = p0 + 32 > k ? (int)(k - p0) : 32;
                for (int j = a; j < b; j++) {
                    const uint32_t cd = code[p0 + j];
                    hi |= ((cd >> 1) & 1u) << j;
                    lo |= (cd & 1u) << j;
                    va |= (cd >> 2) << j;
                }
                uint32_t *dst = planes32 + (hb >> 1) * 6ull + (hb & 1ull);                 // [hi.lo32, hi.hi32, lo.lo32, lo.hi32, va.., va..]
                if (a == 0 && b == 32) { dst[0] = hi; dst[2] = lo; dst[4] = va; }
                else {
                    if (hi) atomicOr(dst + 0, hi);
                    if (lo) atomicOr(dst + 2, lo);
                    if (va) atomicOr(dst + 4, va);
                }
            }
            return;
        }
        __syncthreads();
    }
    tile_kept(txt, n, S, t0, sm, kept, err, &sc);
    // each thread owns 16 consecutive bytes; block-wide exclusive scan of kept counts
    int c = 0;
    const int per = FT / FTHREADS;      // 16
    for (int j = 0; j < per; j++) c += kept[tid * per + j];
    int inc = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    if (lane == 31) wtot[warp] = inc;
    __syncthreads();
    if (tid == 0) {
        int s = 0;
        for (int w = 0; w < FTHREADS / 32; w++) { int x = wtot[w]; wtot[w] = s; s += x; }
        wtot[FTHREADS / 32] = s;
    }
    __syncthreads();
    int pos = wtot[warp] + inc - c;
    const uint64_t tbase = tile_excl[blockIdx.x];
    for (int j = 0; j < per; j++) {
        int i = tid * per + j;
        if (!kept[i]) continue;
        uint32_t ch = sm[i];
        uint32_t cd = ch == 'A' ? 4u : ch == 'C' ? 5u : ch == 'G' ? 6u : ch == 'T' ? 7u : 0u;   // valid<<2 | code
        unsigned long long dummy;
        int r = span_find(S, sc, t0 + i, &dummy);
        code[pos] = (uint8_t)cd;
        opos[pos] = (unsigned long long)(blk_off[r] * 64ull + (tbase + (uint64_t)pos - pbeg[r]));
        pos++;
    }
    __syncthreads();
    const int k = wtot[FTHREADS / 32];
    // one compacted byte per lane: lanes that fall into the same output 64-base block
    // (__match_any_sync) OR their bits together with warp reductions and the group's first
    // lane issues the atomics.  Blocks are shared with neighbouring warps / tiles, hence
    // atomicOr into zero-initialised planes.
    for (int base = 0; base < k; base += FTHREADS) {
        int i = base + tid;
        bool act = i < k;
        unsigned long long o = act ? opos[i] : ~0ull;
        unsigned long long blk = o >> 6;
        uint32_t cd = act ? code[i] : 0u;
        int b = (int)(o & 63);
        unsigned grp = __match_any_sync(0xffffffffu, blk);
        uint32_t blo = b < 32 ? (1u << b) : 0u, bhi = b >= 32 ? (1u << (b - 32)) : 0u;
        uint32_t h0 = __reduce_or_sync(grp, (cd & 2u) ? blo : 0u), h1 = __reduce_or_sync(grp, (cd & 2u) ? bhi : 0u);
        uint32_t l0 = __reduce_or_sync(grp, (cd & 1u) ? blo : 0u), l1 = __reduce_or_sync(grp, (cd & 1u) ? bhi : 0u);
        uint32_t v0 = __reduce_or_sync(grp, (cd & 4u) ? blo : 0u), v1 = __reduce_or_sync(grp, (cd & 4u) ? bhi : 0u);
        if (act && lane == __ffs(grp) - 1) {
            unsigned long long hi = ((unsigned long long)h1 << 32) | h0, lo = ((unsigned long long)l1 << 32) | l0;
            unsigned long long va = ((unsigned long long)v1 << 32) | v0;
            if (hi) atomicOr(&planes[blk * 3 + 0], hi);
            if (lo) atomicOr(&planes[blk * 3 + 1], lo);
            if (va) atomicOr(&planes[blk * 3 + 2], va);
        }
    }
}

// ---- 5. count --------------------------------------------------------------
constexpr int MAXPAT = 64;
struct Patterns {
    int npat;
    int len[MAXPAT];
    unsigned long long hi[MAXPAT], lo[MAXPAT];   // bit j = plane bit of pattern char j
    int border[MAXPAT];                           // 1: pattern can overlap itself
};

// occurrence mask of pattern q for the 64 start positions of block b
__device__ __forceinline__ unsigned long long occ_mask(const Patterns &P, int q, unsigned long long h0, unsigned long long l0,
                                                       unsigned long long v0, unsigned long long h1, unsigned long long l1,
                                                       unsigned long long v1)
{
    unsigned long long m = ~0ull;
    const int L = P.len[q];
    for (int j = 0; j < L; j++) {
        unsigned long long hs = j ? ((h0 >> j) | (h1 << (64 - j))) : h0;
        unsigned long long ls = j ? ((l0 >> j) | (l1 << (64 - j))) : l0;
        unsigned long long vs = j ? ((v0 >> j) | (v1 << (64 - j))) : v0;
        unsigned long long ph = ((P.hi[q] >> j) & 1ull) ? ~0ull : 0ull;
        unsigned long long pl = ((P.lo[q] >> j) & 1ull) ? ~0ull : 0ull;
        m &= ~(hs ^ ph) & ~(ls ^ pl) & vs;
    }
    return m;
}

__device__ __forceinline__ int record_of(const uint64_t *blk_off, int nrec, uint64_t b)
{
    int lo = 0, hi = nrec;
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (blk_off[mid] <= b) lo = mid + 1; else hi = mid;
    }
    return lo - 1;
}

__global__ void __launch_bounds__(256) re_count_kernel(const unsigned long long *planes, const uint64_t *blk_off, const uint64_t *seqlen,
                                                       int nrec, uint64_t nblk, Patterns P, unsigned long long *left,
                                                       unsigned long long *right, unsigned long long *occ /* nullable: [nborder][nblk] */)
{
    __shared__ unsigned long long red[2][8];
    __shared__ int rec0;
    uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int r = -1;
    unsigned long long nl = 0, nr = 0;
    if (b < nblk) {
        r = record_of(blk_off, nrec, b);
        uint64_t len = seqlen[r];
        uint64_t base0 = (b - blk_off[r]) * 64ull;          // in-record index of this block's bit 0
        unsigned long long h0 = planes[b * 3], l0 = planes[b * 3 + 1], v0 = planes[b * 3 + 2];
        unsigned long long h1 = 0, l1 = 0, v1 = 0;
        if (b + 1 < blk_off[r + 1]) { h1 = planes[b * 3 + 3]; l1 = planes[b * 3 + 4]; v1 = planes[b * 3 + 5]; }
        // left half: start < len / 2  (RE_sites.py:71, integer division)
        uint64_t half = len / 2;
        unsigned long long lmask;
        if (half <= base0) lmask = 0;
        else if (half - base0 >= 64) lmask = ~0ull;
        else lmask = (1ull << (half - base0)) - 1ull;
        int bi = 0;
        for (int q = 0; q < P.npat; q++) {
            unsigned long long m = occ_mask(P, q, h0, l0, v0, h1, l1, v1);
            if (P.border[q]) {
                occ[(uint64_t)bi * nblk + b] = m;
                bi++;
            } else {
                nl += __popcll(m & lmask);
                nr += __popcll(m & ~lmask);
            }
        }
    }
    // one atomic pair per CTA when the whole CTA lies in one record (the common case)
    if (threadIdx.x == 0) rec0 = r;
    __syncthreads();
    int same = __syncthreads_and(r == rec0 || r < 0);
    if (same && rec0 >= 0) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            nl += __shfl_xor_sync(0xffffffffu, nl, d);
            nr += __shfl_xor_sync(0xffffffffu, nr, d);
        }
        if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = nl; red[1][threadIdx.x >> 5] = nr; }
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned long long a = 0, c = 0;
            for (int w = 0; w < 8; w++) { a += red[0][w]; c += red[1][w]; }
            if (a) atomicAdd(&left[rec0], a);
            if (c) atomicAdd(&right[rec0], c);
        }
    } else if (r >= 0) {
        if (nl) atomicAdd(&left[r], nl);
        if (nr) atomicAdd(&right[r], nr);
    }
}

// first occurrence in [a, b) of a record's occurrence bits (b - a <= 64), or ~0
__device__ __forceinline__ uint64_t occ_first(const unsigned long long *orec, uint64_t nb, uint64_t a, uint64_t b)
{
    if (a >= b) return ~0ull;
    uint64_t w0 = a >> 6;
    unsigned long long x = w0 < nb ? orec[w0] >> (a & 63) : 0ull;
    if ((a & 63) && w0 + 1 < nb) x |= orec[w0 + 1] << (64 - (a & 63));
    uint64_t span = b - a;
    if (span < 64) x &= (1ull << span) - 1ull;
    return x ? a + (uint64_t)(__ffsll((long long)x) - 1) : ~0ull;
}

// Leftmost non-overlapping selection for a self-overlapping pattern of length L
// (re.finditer resumes the search at the end of a match, RE_sites.py:66).
// An occurrence with no other occurrence in the L-1 bases before it starts a
// chain and is always selected; one thread walks each chain: select s, jump to
// the first occurrence >= s + L, stop when that one starts a chain of its own
// (or lies >= 2L away, which makes it a chain start).
__global__ void re_chain_kernel(const unsigned long long *occ, const uint64_t *blk_off, const uint64_t *seqlen, int nrec,
                                uint64_t nblk, int L, unsigned long long *left, unsigned long long *right)
{
    uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nblk) return;
    unsigned long long m = occ[b];
    if (!m) return;
    int r = record_of(blk_off, nrec, b);
    const unsigned long long *orec = occ + blk_off[r];
    uint64_t nb = blk_off[r + 1] - blk_off[r];
    uint64_t half = seqlen[r] / 2;
    unsigned long long nl = 0, nr = 0;
    while (m) {
        int bit = __ffsll((long long)m) - 1;
        m &= m - 1;
        uint64_t p = (b - blk_off[r]) * 64ull + (uint64_t)bit;
        uint64_t lo = p >= (uint64_t)(L - 1) ? p - (uint64_t)(L - 1) : 0;
        if (occ_first(orec, nb, lo, p) != ~0ull) continue;          // not a chain start
        uint64_t s = p;
        for (;;) {
            if (s < half) nl++; else nr++;
            uint64_t nx = occ_first(orec, nb, s + (uint64_t)L, s + 2ull * (uint64_t)L);
            if (nx == ~0ull) break;
            uint64_t lo2 = nx - (uint64_t)(L - 1);
            if (occ_first(orec, nb, lo2, nx) == ~0ull) break;       // nx starts its own chain
            s = nx;
        }
    }
    if (nl) atomicAdd(&left[r], nl);
    if (nr) atomicAdd(&right[r], nr);
}

inline unsigned nblk_of(size_t n, int t) { return (unsigned)((n + t - 1) / t); }

}   // namespace

// ---------------------------------------------------------------- C ABI
extern "C" int salsa_fasta_headers(const void *d_fasta, size_t nbytes, uint64_t *d_pos, uint32_t *d_count, uint32_t capacity,
                                   void *stream)
{
    if (!d_fasta || !d_pos || !d_count) SALSA_FAIL(SALSA_ERR_ARG, "salsa_fasta_headers: NULL argument");
    cudaStream_t st = as_stream(stream);
    CUDA_TRY(cudaMemsetAsync(d_count, 0, sizeof(uint32_t), st));
    if (nbytes) fasta_headers_kernel<<<148 * 8, 256, 0, st>>>((const uint8_t *)d_fasta, nbytes, (unsigned long long *)d_pos, d_count, capacity);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}

extern "C" size_t salsa_fasta_tiles(size_t nbytes) { return (nbytes + FT - 1) / FT; }

extern "C" int salsa_fasta_count(const void *d_fasta, size_t nbytes, const uint64_t *d_span_beg, const uint64_t *d_span_end,
                                 int32_t nspan, uint64_t *d_tile_cnt, int32_t *d_err, void *stream)
{
    if (!d_fasta || !d_span_beg || !d_span_end || !d_tile_cnt || !d_err) SALSA_FAIL(SALSA_ERR_ARG, "salsa_fasta_count: NULL argument");
    Spans S{(const unsigned long long *)d_span_beg, (const unsigned long long *)d_span_end, nspan};
    size_t nt = salsa_fasta_tiles(nbytes);
    if (nt) fasta_count_kernel<<<(unsigned)nt, FTHREADS, 0, as_stream(stream)>>>((const uint8_t *)d_fasta, nbytes, S, d_tile_cnt, d_err);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}

// d_tile_excl: exclusive scan of the tile counts (size tiles + 1)
extern "C" int salsa_fasta_span_lengths(const void *d_fasta, size_t nbytes, const uint64_t *d_span_beg, const uint64_t *d_span_end,
                                        int32_t nspan, const uint64_t *d_tile_excl, uint64_t *d_pbeg, uint64_t *d_plen,
                                        int32_t *d_err, void *stream)
{
    if (nspan <= 0) return SALSA_OK;
    Spans S{(const unsigned long long *)d_span_beg, (const unsigned long long *)d_span_end, nspan};
    span_prefix_kernel<<<(unsigned)nspan, FTHREADS, 0, as_stream(stream)>>>((const uint8_t *)d_fasta, nbytes, S, d_tile_excl, d_pbeg, d_plen, d_err);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}

extern "C" int salsa_fasta_pack(const void *d_fasta, size_t nbytes, const uint64_t *d_span_beg, const uint64_t *d_span_end,
                                int32_t nspan, const uint64_t *d_tile_excl, const uint64_t *d_pbeg, const uint64_t *d_blk_off,
                                uint64_t *d_planes, int32_t *d_err, void *stream)
{
    if (nspan <= 0 || nbytes == 0) return SALSA_OK;
    Spans S{(const unsigned long long *)d_span_beg, (const unsigned long long *)d_span_end, nspan};
    size_t nt = salsa_fasta_tiles(nbytes);
    fasta_pack_kernel<<<(unsigned)nt, FTHREADS, 0, as_stream(stream)>>>((const uint8_t *)d_fasta, nbytes, S, d_tile_excl, d_pbeg, d_blk_off,
                                                                       (unsigned long long *)d_planes, d_err);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}

// exclusive scan helper shared with links_table.cu
extern "C" int salsa_exclusive_scan_u64(uint64_t *d_vals, size_t n, void *stream);

extern "C" int salsa_re_count(const uint64_t *d_planes, const uint64_t *d_blk_off, const uint64_t *d_seqlen, int32_t nrec,
                              uint64_t nblk, const char *patterns, int32_t npat, uint64_t *d_left, uint64_t *d_right,
                              uint64_t *d_occ_scratch, void *stream)
{
    if (nrec <= 0 || nblk == 0) return SALSA_OK;
    if (!d_planes || !d_blk_off || !d_seqlen || !patterns || !d_left || !d_right || npat < 0 || npat > MAXPAT)
        SALSA_FAIL(SALSA_ERR_ARG, "salsa_re_count: bad argument (at most %d expanded patterns)", MAXPAT);
    Patterns P;
    memset(&P, 0, sizeof P);
    P.npat = npat;
    const char *s = patterns;
    int nborder = 0;
    for (int q = 0; q < npat; q++) {
        int L = (int)strlen(s);
        if (L < 1 || L > 63) SALSA_FAIL(SALSA_ERR_ARG, "pattern %d has length %d (1..63 supported)", q, L);
        P.len[q] = L;
        for (int j = 0; j < L; j++) {
            int cd = s[j] == 'A' ? 0 : s[j] == 'C' ? 1 : s[j] == 'G' ? 2 : s[j] == 'T' ? 3 : -1;
            if (cd < 0) SALSA_FAIL(SALSA_ERR_ARG, "pattern %d contains %c", q, s[j]);
            if (cd & 2) P.hi[q] |= 1ull << j;
            if (cd & 1) P.lo[q] |= 1ull << j;
        }
        // self-overlap border: a proper prefix equal to a suffix
        for (int k = 1; k < L; k++)
            if (memcmp(s, s + L - k, (size_t)k) == 0) P.border[q] = 1;
        nborder += P.border[q];
        s += L + 1;
    }
    if (nborder && !d_occ_scratch) SALSA_FAIL(SALSA_ERR_ARG, "salsa_re_count: %d self-overlapping pattern(s) need d_occ_scratch", nborder);
    cudaStream_t st = as_stream(stream);
    re_count_kernel<<<nblk_of(nblk, 256), 256, 0, st>>>((const unsigned long long *)d_planes, d_blk_off, d_seqlen, nrec, nblk, P,
                                                         (unsigned long long *)d_left, (unsigned long long *)d_right,
                                                         (unsigned long long *)d_occ_scratch);
    int bi = 0;
    for (int q = 0; q < npat; q++) {
        if (!P.border[q]) continue;
        re_chain_kernel<<<nblk_of(nblk, 256), 256, 0, st>>>((const unsigned long long *)d_occ_scratch + (uint64_t)bi * nblk, d_blk_off,
                                                             d_seqlen, nrec, nblk, P.len[q], (unsigned long long *)d_left,
                                                             (unsigned long long *)d_right);
        bi++;
    }
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}


// ---- per-scaffold RE counts for iterations >= 2 ---------------------------------------------------------------
// Replaces the proration loop of layout_unitigs.py:736-814 / refactor_breaks.py:63-131: a scaffold's contigs are laid end
// to end; RE sites of contigs left of the midpoint (Python-2 integer length/2) count left, right of it right, and the
// contig(s) that touch the midpoint are split in proportion -- float(left+right)/len * part, summed in PATH ORDER in
// double precision (ints and floats mix exactly as CPython's int/float additions do while sums stay below 2^53) and
// truncated by int().  A single-contig scaffold keeps its contig's (left, right), swapped when it is reversed.
namespace {
__global__ void scaffold_re_kernel(const int64_t *path_off, const int32_t *path_contig, const uint8_t *first_is_b, const int64_t *clen,
                                   const int64_t *re_left, const int64_t *re_right, int32_t nscaf, int64_t *out_left, int64_t *out_right)
{
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= nscaf) return;
    const int64_t b = path_off[s], e = path_off[s + 1];
    if (e - b == 1) {
        const int32_t c = path_contig[b];
        const int64_t l = re_left[c], r = re_right[c];
        out_left[s] = first_is_b[s] ? l : r;
        out_right[s] = first_is_b[s] ? r : l;
        return;
    }
    int64_t total = 0;
    for (int64_t i = b; i < e; i++) total += clen[path_contig[i]];
    const int64_t mid = total / 2;                       // lengths are >= 0: floor == Python-2 '/'
    double sl = 0.0, sr = 0.0;
    int64_t off = 0;
    for (int64_t i = b; i < e; i++) {
        const int32_t c = path_contig[i];
        const int64_t len = clen[c], lr = re_left[c] + re_right[c];
        const int64_t start = off, end = off + len;
        if (end < mid) sl = __dadd_rn(sl, (double)lr);
        if (start > mid) sr = __dadd_rn(sr, (double)lr);
        if (start <= mid && end >= mid) {
            const double per = __ddiv_rn((double)lr, (double)len);
            sl = __dadd_rn(sl, __dmul_rn(per, (double)(mid - start)));
            sr = __dadd_rn(sr, __dmul_rn(per, (double)(end - mid)));
        }
        off = end;
    }
    out_left[s] = (int64_t)sl;                            // int(): truncation toward zero
    out_right[s] = (int64_t)sr;
}
}   // namespace

extern "C" int salsa_scaffold_re(const int64_t *d_path_off, const int32_t *d_path_contig, const uint8_t *d_first_is_b,
                                 const int64_t *d_contig_len, const int64_t *d_re_left, const int64_t *d_re_right, int32_t nscaf,
                                 int64_t *d_out_left, int64_t *d_out_right, void *stream)
{
    if (nscaf == 0) return SALSA_OK;
    if (nscaf < 0 || !d_path_off || !d_path_contig || !d_first_is_b || !d_contig_len || !d_re_left || !d_re_right || !d_out_left ||
        !d_out_right)
        SALSA_FAIL(SALSA_ERR_ARG, "salsa_scaffold_re: bad argument");
    scaffold_re_kernel<<<(unsigned)((nscaf + 127) / 128), 128, 0, as_stream(stream)>>>(d_path_off, d_path_contig, d_first_is_b,
                                                                                     d_contig_len, d_re_left, d_re_right, nscaf,
                                                                                     d_out_left, d_out_right);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}
