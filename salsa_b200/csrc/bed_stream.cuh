// bed_stream.cuh -- the streaming BED scanner of the link histogram (make_links.py:60-146).
//
// Where bed_scan.cuh cuts the text into tiles that CTAs pick up one by one, this kernel is
// PERSISTENT and WARP-AUTONOMOUS: the buffer is cut into 256 KiB chunks that the resident warps
// draw, in file order, from one atomic counter, and every warp walks its chunk from start to end
// through a private 8 KiB ring in shared memory that it fills itself with cp.async (LDGSTS.128,
// 16 bytes per lane, L1 bypassed; 2 KiB groups, up to three in flight).  There is no block-wide
// barrier, no halo that is classified twice, no per-tile prologue, and the "previous line" simply
// travels along the chunk.  (A first version staged the ring with one 1-D TMA bulk copy + mbarrier
// per warp and stage: the elected-lane issue sequence cost more instructions than 32 lanes x 16 B.)
//
// Per warp, three interleaved activities:
//   CLASSIFY  2 KiB per iteration (64 B per lane): SWAR byte tests + dp4a bit packing give the
//             "special" bitmap (one bit per ring byte: <= 0x20 or '/', or the cheaper superset
//             <= 0x2F when a sample of the text allows it) and, from the same flags, the newline
//             mask; line starts are ranked by ballot and appended to a small ring of line starts.
//   PHASE 1   32 lines per step, one lane per line: the four field boundaries from the special
//             bitmap, separator bytes checked, the contig name as a masked 16-byte window, the
//             last four bytes of read.split('/')[0].  A pair of plain lines whose names are equal
//             or whose read names provably differ is dropped here (exactly the pairs the
//             reference neither counts nor dies on); everything else is queued (16 bytes).
//   PHASE 2   16 queued pairs at a time, two lanes per pair (one line each), the lines re-read
//             from global memory (L2): name -> id through the device hash table, start/end by
//             dp4a digit conversion, full read-name compare, then the reference's checks in the
//             reference's order and one RED into the table.  A line that is not of the plain form
//             is parsed by the byte-loop parser straight from global memory (str.split() in
//             full): identical results, only slower.
#pragma once
#include "bed_scan.cuh"

namespace bedstream {
using namespace bedscan;

#ifndef SALSA_STREAM_RING
#define SALSA_STREAM_RING 8192
#endif
#ifndef SALSA_STREAM_WARPS
#define SALSA_STREAM_WARPS 20
#endif
constexpr int RING = SALSA_STREAM_RING;     // bytes of text a warp keeps in shared memory
constexpr int GRP = 2048;                   // bytes per cp.async group = one classification iteration (64 bytes per lane)
constexpr int NGR = RING / GRP;
#ifndef SALSA_STREAM_MIRROR
#define SALSA_STREAM_MIRROR 256
#endif
#ifndef SALSA_STREAM_LCAP
#define SALSA_STREAM_LCAP 256
#endif
constexpr int MIRROR = SALSA_STREAM_MIRROR;  // the ring's first bytes repeated behind its end: reads never wrap (>= 128: phase 1 reads 96 bitmap bits and ~72 bytes from a line's start)
constexpr int FRONT = 16;                   // addressable bytes before the ring (masked-off window bytes)
constexpr int OVER = 160;                   // phase 1 looks at a line once this much behind its start is classified
constexpr int LCAP = SALSA_STREAM_LCAP;      // ring of line starts (entries; >= 128: an iteration adds up to 64 without asking)
constexpr int QCAP = 64;                    // ring of queued pairs (entries of 16 bytes)
constexpr int QPAIRS = 16;                  // pairs per phase-2 batch (two lanes each)
constexpr int WARPS = SALSA_STREAM_WARPS;   // warps per CTA (they share nothing but the mask table)
constexpr int SPW = (RING + MIRROR) / 32 + 4;
constexpr uint32_t CHUNK_MIN = 16384;       // shortest chunk worth a warp
#ifndef SALSA_STREAM_CHUNK
#define SALSA_STREAM_CHUNK 262144
#endif
constexpr uint32_t CHUNK_MAX = SALSA_STREAM_CHUNK;   // bytes of text a warp walks through before it takes the next chunk
static_assert(RING % GRP == 0 && NGR >= 4 && NGR <= 8 && (RING & (RING - 1)) == 0, "ring geometry");

constexpr size_t WARP_SMEM = FRONT + RING + MIRROR + SPW * 4 + LCAP * 4 + QCAP * 16;
static_assert(WARP_SMEM % 16 == 0, "per-warp shared memory must keep 16-byte alignment");
constexpr size_t CTA_SMEM = 17 * 16 + WARPS * WARP_SMEM;

// ---- what phase 2 knows about one line ----
// info bits: 0-2 nf, 3 numbad, 4 toolong, 5 has_len, 6 has_re, 7 len == 0, 8 end is 'E', 9 parsed from global memory
struct LineInfo {
    int id;
    uint32_t info;
    unsigned long long noff, roff;      // abase coordinates of the contig name and of the read token
    uint32_t nlen, rlen;                // name length, length of read.split('/')[0]
};

__device__ __forceinline__ uint32_t gbyte(const ScanArgs &A, unsigned long long q, unsigned long long lim)
{
    return q < lim ? (uint32_t)__ldg(A.abase + q) : 10u;
}

// str.split() on the line that starts at abase coordinate st, read from global memory (make_links.py:62,70).
// The first four fields must end within SALSA_MAX_LINE bytes of the line's start (library limit).
static __device__ __noinline__ void parse_global(const ScanArgs &A, unsigned long long st, LineInfo &L, long long &sum)
{
    unsigned long long lim = st + SALSA_MAX_LINE;
    const bool capped = lim < A.dend;
    if (!capped) lim = A.dend;
    unsigned long long p = st;
    L.id = -1; L.info = 1u << 9; L.noff = L.roff = 0; L.nlen = L.rlen = 0;
    sum = 0;
    uint32_t c = gbyte(A, p, lim);
    while (c != '\n' && is_ws(c)) c = gbyte(A, ++p, lim);
    uint32_t nf = 0, bad = 0;
    if (c != '\n') {
        unsigned long long noff = p;
        do { c = gbyte(A, ++p, lim); } while (!is_ws(c));
        uint32_t nlen = (uint32_t)(p - noff);
        L.noff = noff; L.nlen = nlen;
        uint32_t w[4] = {0, 0, 0, 0};
        for (uint32_t i = 0; i < nlen && i < 16; i++) w[i >> 2] |= (uint32_t)__ldg(A.abase + noff + i) << (8 * (i & 3));
        {
            const DevTable &t = A.t;
            uint32_t h = salsa_name_hash(w[0], w[1], w[2], w[3], nlen) & t.mask;
            for (;;) {
                const uint4 *sp = reinterpret_cast<const uint4 *>(&t.slots[h]);
                uint4 k = __ldg(sp), m = __ldg(sp + 1);
                if (m.x == 0xFFFFFFFFu) break;
                if (m.x == nlen && k.x == w[0] && k.y == w[1] && k.z == w[2] && k.w == w[3]) {
                    bool eq = true;
                    for (uint32_t i = 16; i < nlen; i++)
                        if ((uint8_t)__ldg(&t.blob[m.z + i]) != __ldg(A.abase + noff + i)) { eq = false; break; }
                    if (eq) { L.id = (int)m.y; break; }
                }
                h = (h + 1) & t.mask;
            }
        }
        nf = 1;
        long long val[2] = {0, 0};
#pragma unroll
        for (int f = 0; f < 2; f++) {
            if (nf == 1u + f) {
                while (c != '\n' && is_ws(c)) c = gbyte(A, ++p, lim);
                if (c != '\n') {
                    nf++;
                    bool neg = (c == '-');
                    if (c == '-' || c == '+') c = gbyte(A, ++p, lim);
                    unsigned long long v = 0;
                    int nd = 0;
                    uint32_t d = c - '0';
                    while (d <= 9u) { v = v * 10ull + d; nd++; c = gbyte(A, ++p, lim); d = c - '0'; }
                    if (nd == 0 || nd > 18 || !is_ws(c)) bad = 8u;
                    while (!is_ws(c)) c = gbyte(A, ++p, lim);
                    val[f] = neg ? -(long long)v : (long long)v;
                }
            }
        }
        sum = val[0] + val[1];
        if (nf == 3) {
            while (c != '\n' && is_ws(c)) c = gbyte(A, ++p, lim);
            if (c != '\n') {
                nf = 4;
                unsigned long long roff = p;
                while (c != '/' && !is_ws(c)) c = gbyte(A, ++p, lim);      // read.split('/')[0]
                L.roff = roff; L.rlen = (uint32_t)(p - roff);
            }
        }
    }
    L.info |= nf | bad | ((capped && p >= lim) ? 16u : 0u);
}

__device__ __forceinline__ bool gbytes_equal(const ScanArgs &A, unsigned long long a, unsigned long long b, uint32_t n)
{
    for (uint32_t i = 0; i < n; i++)
        if (__ldg(A.abase + a + i) != __ldg(A.abase + b + i)) return false;
    return true;
}

// ---- phase 1: the field layout of one line ----
// P = p1 | p2 << 8 | p3 << 16 | rl << 24 (offsets of the first three separators from the line's start, length of the
// read prefix); 0 = the line is not provably of the plain form  name SEP digits SEP digits SEP read[/...]  with
// single tab/space separators, a name of at most 31 and a read prefix of at most ~35 bytes.  w0..w3 = the name's first
// 16 bytes, zero padded (plain names hold no byte <= 0x20, so equal windows <=> equal names up to 16 bytes);
// d = the last four bytes of the read prefix (whatever precedes it when it is shorter).
struct Shp {
    uint32_t w0, w1, w2, w3, P, d;
};

__device__ __forceinline__ Shp shape_of(const uint8_t *s, const uint32_t *spw, const uint4 *LM, uint32_t p, uint32_t room)
{
    Shp h;
    const uint32_t i = p >> 5, sh = p & 31u;
    const uint32_t a0 = spw[i], a1 = spw[i + 1], a2 = spw[i + 2];
    // the 64 bits from the line's start on, bit-reversed: the next special byte is then the leading one (one FLO), and
    // shifting left drops it.  A distance of 32 = none in reach.
    uint32_t rh = __brev(__funnelshift_r(a0, a1, sh)), rl = __brev(__funnelshift_r(a1, a2, sh));
    const uint32_t d1 = __clz(rh);
    rh = __funnelshift_lc(rl, rh, d1 + 1u); rl = __funnelshift_lc(0u, rl, d1 + 1u);
    const uint32_t d2 = __clz(rh);
    rh = __funnelshift_lc(rl, rh, d2 + 1u); rl = __funnelshift_lc(0u, rl, d2 + 1u);
    const uint32_t d3 = __clz(rh);
    rh = __funnelshift_lc(rl, rh, d3 + 1u); rl = __funnelshift_lc(0u, rl, d3 + 1u);
    const uint32_t d4 = rh ? __clz(rh) : 32u + __clz(rl);
    const uint32_t p1 = d1, p2 = p1 + 1u + d2, p3 = p2 + 1u + d3, p4 = p3 + 1u + d4;
    const uint32_t c1 = s[p + p1], c2 = s[p + p2], c3 = s[p + p3], c4 = s[p + p4];
    // no short-circuit: every lane runs the same few instructions
    const bool ok = ((((d1 | d2 | d3) & 32u) | (d4 & 64u)) == 0u) & ((d1 * d2) * (d3 * d4) != 0u) & (p4 < room) &
                    ((c1 == 9u) | (c1 == 32u)) & ((c2 == 9u) | (c2 == 32u)) & ((c3 == 9u) | (c3 == 32u)) &
                    ((c4 == 47u) | (c4 == 9u) | (c4 == 32u) | (c4 == 10u));
    const uint32_t *s32 = reinterpret_cast<const uint32_t *>(s);
    const uint32_t a = p >> 2, shn = (p & 3u) * 8u;
    const uint32_t x0 = s32[a], x1 = s32[a + 1], x2 = s32[a + 2], x3 = s32[a + 3], x4 = s32[a + 4];
    const uint4 mk = LM[p1 > 16u ? 16u : p1];
    h.w0 = __funnelshift_r(x0, x1, shn) & mk.x;
    h.w1 = __funnelshift_r(x1, x2, shn) & mk.y;
    h.w2 = __funnelshift_r(x2, x3, shn) & mk.z;
    h.w3 = __funnelshift_r(x3, x4, shn) & mk.w;
    const int q = (int)(p + p4) - 4;                     // p4 >= 7 on a plain line; >= -1 always (FRONT pad)
    const uint32_t y0 = s32[q >> 2], y1 = s32[(q >> 2) + 1];
    h.d = __funnelshift_r(y0, y1, ((uint32_t)q & 3u) * 8u);
    h.P = ok ? (p1 | (p2 << 8) | (p3 << 16) | (d4 << 24)) : 0u;
    return h;
}

// the n (1..12) decimal digits that END just before ring offset b -> value; false if a byte is not a digit
__device__ __forceinline__ bool ring_digits12(const uint8_t *s, const uint4 *LM, uint32_t b, uint32_t n, long long &val)
{
    const uint32_t *s32 = reinterpret_cast<const uint32_t *>(s);
    const int q = (int)b - 12;                           // >= -9: inside the FRONT pad
    const int a = q >> 2;
    const uint32_t sh = ((uint32_t)q & 3u) * 8u;
    const uint32_t x0 = s32[a], x1 = s32[a + 1], x2 = s32[a + 2], x3 = s32[a + 3];
    const uint4 mk = LM[12u - n];                        // the low 12 - n bytes of the window precede the field
    const uint32_t u0 = (__funnelshift_r(x0, x1, sh) ^ 0x30303030u) & ~mk.x;
    const uint32_t u1 = (__funnelshift_r(x1, x2, sh) ^ 0x30303030u) & ~mk.y;
    const uint32_t u2 = (__funnelshift_r(x2, x3, sh) ^ 0x30303030u) & ~mk.z;
    const uint32_t bad = ((((u0 & 0x7F7F7F7Fu) + 0x76767676u) | u0) | (((u1 & 0x7F7F7F7Fu) + 0x76767676u) | u1) |
                          (((u2 & 0x7F7F7F7Fu) + 0x76767676u) | u2)) & 0x80808080u;
    const uint32_t v0 = __dp4a(u0, 0x0000010Au, 0u) * 100u + __dp4a(u0, 0x010A0000u, 0u);
    const uint32_t v1 = __dp4a(u1, 0x0000010Au, 0u) * 100u + __dp4a(u1, 0x010A0000u, 0u);
    const uint32_t v2 = __dp4a(u2, 0x0000010Au, 0u) * 100u + __dp4a(u2, 0x010A0000u, 0u);
    val = (long long)((unsigned long long)(v0 * 10000u + v1) * 10000ull + v2);
    return bad == 0u;
}

// ---- which "special byte" test the classification runs ----
// Exact: byte <= 0x20 or byte == '/' (7 instructions per 4-byte word).  Cheap: byte <= 0x2F (3 per word) -- a superset:
// a line whose first four fields hold one of the 14 characters 0x21..0x2E ('.', '-', '#', ...) then merely fails the
// separator checks of phase 1 and goes through the full parser, so results are identical either way and only speed is
// at stake.  The sample kernel looks at three windows of the text and picks the cheap test when such characters are no
// more frequent than a strand column would make them (one per line).
__device__ __forceinline__ uint32_t le_bytes_0x2f(uint32_t w)
{
    uint32_t t = (w & 0x7F7F7F7Fu) + 0x50505050u;     // bit 7 set iff (b & 0x7f) > 0x2f
    return ~(t | w) & 0x80808080u;
}
__device__ __forceinline__ uint32_t cheap16_of(const uint4 &v)
{
    return pack16(le_bytes_0x2f(v.x), le_bytes_0x2f(v.y), le_bytes_0x2f(v.z), le_bytes_0x2f(v.w));
}

// special flags (0x80 per special byte) of one word -> newline flags: a special byte has bit 7 clear, so it is '\n' iff the low
// seven bits of (byte ^ 0x0A) are zero; (y | 0x80) - 1 clears bit 7 exactly then, and no byte borrows from its neighbour.
// Three instructions per word where the stand-alone equality test takes four.
__device__ __forceinline__ uint32_t nl_of_special(uint32_t w, uint32_t spf)
{
    const uint32_t u = ((w ^ 0x0A0A0A0Au) | 0x80808080u) - 0x01010101u;
    return ~u & spf;
}
template <bool CHEAP>
__device__ __forceinline__ void masks16(const uint4 &v, uint32_t &sp, uint32_t &nl)
{
    const uint32_t f0 = CHEAP ? le_bytes_0x2f(v.x) : (le_bytes_0x20(v.x) | eq_bytes(v.x, 0x2F2F2F2Fu));
    const uint32_t f1 = CHEAP ? le_bytes_0x2f(v.y) : (le_bytes_0x20(v.y) | eq_bytes(v.y, 0x2F2F2F2Fu));
    const uint32_t f2 = CHEAP ? le_bytes_0x2f(v.z) : (le_bytes_0x20(v.z) | eq_bytes(v.z, 0x2F2F2F2Fu));
    const uint32_t f3 = CHEAP ? le_bytes_0x2f(v.w) : (le_bytes_0x20(v.w) | eq_bytes(v.w, 0x2F2F2F2Fu));
    sp = pack16(f0, f1, f2, f3);
    nl = pack16(nl_of_special(v.x, f0), nl_of_special(v.y, f1), nl_of_special(v.z, f2), nl_of_special(v.w, f3));
}

struct CheapTest { static constexpr bool value = true; };
struct ExactTest { static constexpr bool value = false; };

constexpr int SAMPLE_WIN = 16384;      // bytes per window; three windows: start, middle, end of the text
__global__ void __launch_bounds__(256) bed_sample_kernel(ScanArgs A, int names_punct, int *d_mode)
{
    __shared__ unsigned int cnt[2];            // newlines, bytes in 0x21..0x2E
    if (threadIdx.x < 2) cnt[threadIdx.x] = 0;
    __syncthreads();
    const unsigned long long n16 = A.dend & ~15ull;          // whole 16-byte granules of [abase, abase + dend)
    unsigned int nl = 0, pu = 0;
    for (int w = 0; w < 3; w++) {
        unsigned long long b = w == 0 ? 0ull : (w == 1 ? (n16 / 2) & ~15ull : (n16 > (unsigned long long)SAMPLE_WIN ? n16 - SAMPLE_WIN : 0ull));
        if (w > 0 && n16 <= (unsigned long long)SAMPLE_WIN) break;
        for (unsigned long long i = b + 16ull * threadIdx.x; i < b + SAMPLE_WIN && i < n16; i += 16ull * 256) {
            const uint4 v = __ldg(reinterpret_cast<const uint4 *>(A.abase + i));
            const uint32_t ws[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 4; k++) {
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const uint32_t c = (ws[k] >> (8 * q)) & 255u;
                    if (i + 4 * k + q < A.mis) continue;      // bytes before the caller's pointer are not text
                    nl += c == 10u;
                    pu += (c - 0x21u) <= 0x0Du;
                }
            }
        }
    }
    atomicAdd(&cnt[0], nl);
    atomicAdd(&cnt[1], pu);
    __syncthreads();
    if (threadIdx.x == 0) *d_mode = (!names_punct && cnt[1] <= cnt[0] + cnt[0] / 16u + 2u) ? 1 : 0;
}

// cp.async.wait_group takes an immediate: wait until at most `ahead` of this thread's groups are still in flight
__device__ __forceinline__ void wait_groups(uint32_t ahead)
{
    if (ahead >= 6u) asm volatile("cp.async.wait_group 6;" ::: "memory");
    else if (ahead == 5u) asm volatile("cp.async.wait_group 5;" ::: "memory");
    else if (ahead == 4u) asm volatile("cp.async.wait_group 4;" ::: "memory");
    else if (ahead == 3u) asm volatile("cp.async.wait_group 3;" ::: "memory");
    else if (ahead == 2u) asm volatile("cp.async.wait_group 2;" ::: "memory");
    else if (ahead == 1u) asm volatile("cp.async.wait_group 1;" ::: "memory");
    else asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// one aligned 32-bit word of the text in global memory (L2: the bytes went through it moments ago)
__device__ __forceinline__ const uint32_t *gwords(const uint8_t *g, unsigned long long off)
{
    return reinterpret_cast<const uint32_t *>(g + (off & ~3ull));
}

// the n (1..12) decimal digits that END just before global offset b (b >= 16)
__device__ __forceinline__ bool global_digits12(const uint8_t *g, const uint4 *LM, unsigned long long b, uint32_t n, long long &val)
{
    const unsigned long long q = b - 12ull;
    const uint32_t sh = ((uint32_t)q & 3u) * 8u;
    const uint32_t *gp = gwords(g, q);
    const uint32_t x0 = __ldg(gp), x1 = __ldg(gp + 1), x2 = __ldg(gp + 2), x3 = __ldg(gp + 3);
    const uint4 mk = LM[12u - n];                        // the low 12 - n bytes of the window precede the field
    const uint32_t u0 = (__funnelshift_r(x0, x1, sh) ^ 0x30303030u) & ~mk.x;
    const uint32_t u1 = (__funnelshift_r(x1, x2, sh) ^ 0x30303030u) & ~mk.y;
    const uint32_t u2 = (__funnelshift_r(x2, x3, sh) ^ 0x30303030u) & ~mk.z;
    const uint32_t bad = ((((u0 & 0x7F7F7F7Fu) + 0x76767676u) | u0) | (((u1 & 0x7F7F7F7Fu) + 0x76767676u) | u1) |
                          (((u2 & 0x7F7F7F7Fu) + 0x76767676u) | u2)) & 0x80808080u;
    const uint32_t v0 = __dp4a(u0, 0x0000010Au, 0u) * 100u + __dp4a(u0, 0x010A0000u, 0u);
    const uint32_t v1 = __dp4a(u1, 0x0000010Au, 0u) * 100u + __dp4a(u1, 0x010A0000u, 0u);
    const uint32_t v2 = __dp4a(u2, 0x0000010Au, 0u) * 100u + __dp4a(u2, 0x010A0000u, 0u);
    val = (long long)((unsigned long long)(v0 * 10000u + v1) * 10000ull + v2);
    return bad == 0u;
}

// name (first 16 bytes in w0..w3, the rest in global memory at noff + 16 ...) -> contig id or -1
__device__ __forceinline__ int lookup_global(const DevTable &t, const uint8_t *g, uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3,
                                             unsigned long long noff, uint32_t nlen, uint32_t h)
{
    for (;;) {
        const uint4 *sp = reinterpret_cast<const uint4 *>(&t.slots[h]);
        const uint4 k = __ldg(sp), m = __ldg(sp + 1);     // words; len, id, blob_off, pad
        if (m.x == 0xFFFFFFFFu) return -1;
        if (m.x == nlen && k.x == w0 && k.y == w1 && k.z == w2 && k.w == w3) {
            bool eq = true;
            for (uint32_t i = 16; i < nlen; i++)
                if ((uint8_t)__ldg(&t.blob[m.z + i]) != __ldg(g + noff + i)) { eq = false; break; }
            if (eq) return (int)m.y;
        }
        h = (h + 1) & t.mask;
    }
}

template <class Emit>
__global__ void __launch_bounds__(WARPS * 32) bed_stream_kernel(typename Emit::Args EA, ScanArgs A, uint32_t unit_bytes, uint32_t nunits, uint32_t take_big, uint32_t tail_units, int *d_mode)
{
    extern __shared__ __align__(128) uint8_t smem[];
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // ptxas would rather re-read %tid and redo the address arithmetic inside the loops than keep these in registers
    asm volatile("" : "+r"(lane), "+r"(warp));
    uint4 *LM = reinterpret_cast<uint4 *>(smem);          // LM[e]: the low e bytes of a 16-byte window
    if (threadIdx.x < 17) {
        uint32_t m[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            int nbv = (int)threadIdx.x - 4 * j;
            m[j] = nbv >= 4 ? 0xFFFFFFFFu : (nbv <= 0 ? 0u : ((1u << (8 * nbv)) - 1u));
        }
        LM[threadIdx.x] = make_uint4(m[0], m[1], m[2], m[3]);
    }
    uint8_t *base = smem + 17 * 16 + (size_t)warp * WARP_SMEM;
    asm volatile("" : "+l"(base));
    __builtin_assume(__isShared(base));
    uint8_t *s = base + FRONT;                                                  // the ring
    uint32_t *spw = reinterpret_cast<uint32_t *>(s + RING + MIRROR);            // special bitmap, bit per ring byte
    uint32_t *lines = spw + SPW;                                                // ring of line starts (stream offsets)
    uint4 *queue = reinterpret_cast<uint4 *>(lines + LCAP);                     // (cur start, prev start, P cur, P prev)
    if (lane < 4) reinterpret_cast<uint32_t *>(base)[lane] = 0u;
    __syncthreads();                                                            // the mask table; the only block-wide barrier

    uint32_t nlines = 0, npairs = 0;
    const unsigned ltmask = (1u << lane) - 1u;
    const bool cheap = *d_mode != 0;             // which special-byte test (same for the whole launch)
    const unsigned long long dend = A.dend, dend16 = (dend + 15ull) & ~15ull;
    // Chunks are handed out in file order through one counter (d_mode[1], zeroed before the launch): at any time the warps of
    // the whole GPU work side by side in one window of the text (a few hundred MB), which keeps the TLB and the DRAM pages they
    // touch few, and a warp that drew chunks with many candidate pairs simply takes fewer of them (dealing them round robin
    // left the last of 61 rounds 40 % empty).
    for (;;) {
    // a chunk = take_big units while plenty of text is left, one unit over the last tail_units: the warps then finish
    // within a quarter of a chunk's time of one another (any cut at unit boundaries is a valid set of chunks)
    uint32_t ck = 0, take = 1;
    if (lane == 0) {
        const uint32_t seen = (uint32_t)*((volatile int *)(d_mode + 1));
        take = (seen < nunits && nunits - seen > tail_units) ? take_big : 1u;
        ck = (uint32_t)atomicAdd(d_mode + 1, (int)take);
    }
    ck = __shfl_sync(0xffffffffu, ck, 0);
    take = __shfl_sync(0xffffffffu, take, 0);
    if (ck >= nunits) break;
    // ---- the chunk, in abase coordinates ----
    const unsigned long long S0 = (unsigned long long)ck * unit_bytes;
    unsigned long long S1 = S0 + (unsigned long long)take * unit_bytes;
    if (S1 > dend || ck + take >= nunits) S1 = dend;
    unsigned long long lo = S0 >= (unsigned long long)HALO ? S0 - HALO : 0ull;            // first byte loaded
    asm volatile("" : "+l"(lo));             // (kept in registers: see lane / warp above)
    unsigned long long hi = S1 + HALO;                                                    // one past the last byte classified
    const bool at_end = hi >= dend16;
    if (at_end) hi = dend16;
    const uint8_t *gsrc = A.abase + lo;
    asm volatile("" : "+l"(gsrc));
    // ---- the same, relative to lo (32 bit) ----
    const uint32_t rel_end = (uint32_t)(hi - lo);
    const uint32_t dend_rel = at_end ? (uint32_t)(dend - lo) : 0xFFFFFF00u;
    const uint32_t lim = at_end ? dend_rel + 1u : rel_end;             // a plain line's fourth separator lies below this
    const uint32_t own_lo = lo == 0 ? A.mis : (uint32_t)(S0 - lo);     // lines that start beyond it are this chunk's
    const uint32_t nl_lo = lo == 0 ? A.mis : 0u;                       // newlines q in [nl_lo, nl_hi) open a line we look at
    const unsigned long long nlh = S1 < dend - 1 ? S1 : dend - 1;
    const uint32_t nl_hi = (uint32_t)(nlh - lo);
    const uint32_t ngr = (rel_end + GRP - 1) / GRP;                     // cp.async groups of this chunk
    const uint32_t wr_end = at_end ? rel_end + 64u : rel_end;          // bitmap words are written up to here (sentinel included)
    uint64_t file_base = A.base_offset - (uint64_t)A.mis + lo;          // file offset of chunk offset 0
    asm volatile("" : "+l"(file_base));
    // phase 2 reads a plain line's words from global memory: keep clear of both ends of the buffer
    const uint32_t g_lo = lo >= 16ull ? 0u : 16u - (uint32_t)lo;
    const uint32_t g_hi = !at_end ? 0xFFFFFF00u : (dend_rel >= 192u ? dend_rel - 192u : 0u);

    // the last group (when the text ends in this chunk) needs room for the sentinel behind it and gets it written
    uint32_t ngr_sent = at_end ? ngr : 0xFFFFFFFFu;
    // these are used inside the loops below: keep them in registers instead of recomputing them from 64-bit values
    uint32_t rel_end_r = rel_end, lim_r = lim, own_lo_r = own_lo, nl_lo_r = nl_lo, nl_hi_r = nl_hi, wr_end_r = wr_end, ngr_r = ngr;
    asm volatile("" : "+r"(rel_end_r), "+r"(lim_r), "+r"(own_lo_r), "+r"(nl_lo_r), "+r"(nl_hi_r), "+r"(wr_end_r), "+r"(ngr_r), "+r"(ngr_sent));
    const uint32_t sring = smem_u32(s) + (uint32_t)lane * 16u;           // this lane's 16 bytes of a cp.async
    const uint8_t *glane = gsrc + lane * 16;
    asm volatile("" : "+l"(glane));

    uint32_t cpos = 0, lh = 0, lt = 0, qh = 0, qn = 0;
    uint32_t gi = 0, gw = 0;             // groups issued / groups that have arrived
    uint32_t over_eff = OVER;            // a line is looked at once this much text behind its start is classified (0 at the end)
    uint32_t next_need = 0;              // the next group can be issued once low has reached this
    Shp hp;                              // the shapes of the previous step (lane 31's is the next step's "previous")
    hp.w0 = hp.w1 = hp.w2 = hp.w3 = hp.P = hp.d = 0;
    uint32_t hp_start = 0;
    uint32_t low = 0;                    // ring bytes from here on are still needed (start of the last line looked at)
    bool have_prev = false;

    if (lo == 0) {                      // the buffer's first line: it only ever plays "previous" (make_links.py:66-68)
        if (lane == 0) lines[0] = A.mis;
        lt = 1;
    }
    __syncwarp();

    // ---- text -> ring: cp.async (LDGSTS, 16 bytes per lane, L1 bypassed), one group per KiB, as many in flight as the ring has room for ----
    auto issue_allowed = [&]() {
        // group g overwrites what group g - NGR held
        while (low >= next_need) {
            const uint32_t goff = gi * GRP;
            const uint32_t d0 = sring + (goff & (RING - 1));
            const uint8_t *src = glane + goff;
            const bool mirror = (goff & (RING - 1)) == 0u && lane < MIRROR / 16;     // the ring's first bytes once more behind its end
            if (gi + 1u < ngr_r) {
#pragma unroll
                for (int j = 0; j < GRP / 512; j++)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d0 + 512u * j), "l"(src + 512 * j) : "memory");
                if (mirror) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d0 + (uint32_t)RING), "l"(src) : "memory");
            } else {
                // the chunk's last group: bytes beyond the text are not read (zero fill)
                const uint32_t r0 = goff + (uint32_t)lane * 16u;
#pragma unroll
                for (int j = 0; j < GRP / 512; j++)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d0 + 512u * j), "l"(src + 512 * j),
                                 "r"(r0 + 512u * j < rel_end_r ? 16u : 0u)
                                 : "memory");
                if (mirror)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d0 + (uint32_t)RING), "l"(src), "r"(r0 < rel_end_r ? 16u : 0u) : "memory");
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
            gi++;
            // what group gi needs: the ring bytes of group gi - NGR free (one group more for the sentinel behind the last one)
            const uint32_t tt = gi + (gi + 1u == ngr_sent ? 2u : 1u);
            next_need = gi >= ngr_r ? 0xFFFFFFFFu : (tt >= (uint32_t)NGR ? (tt - NGR) * GRP : 0u);
        }
    };
    auto wait_for = [&](uint32_t g) {             // group g has been issued: wait for it (and so for all before it)
        wait_groups(gi - (g + 1u));
        __syncwarp();                             // every lane waited for its own copies
        gw = g + 1u;
        if (gw == ngr_sent) {                     // sentinel: the data ends like a line ends
#pragma unroll
            for (int j = 0; j < 2; j++) {
                uint32_t o = (dend_rel + (uint32_t)lane * 2u + j) & (RING - 1);
                s[o] = '\n';
                if (o < (uint32_t)MIRROR) s[RING + o] = '\n';
            }
            __syncwarp();
        }
    };

    // ---- phase 2: nb (<= 16) queued pairs, two lanes per pair ----
    auto flush = [&](uint32_t nb) {
#if defined(SALSA_ABL) && SALSA_ABL == 1          // timing experiments only (tools/sweep_stream.sh): phase 2 dropped
        qh += nb; qn -= nb; return;
#endif
        const uint32_t ei = (uint32_t)lane >> 1, role = (uint32_t)lane & 1u;    // role 0: the pair's current line, 1: its previous line
        const bool act = ei < nb;
        uint4 e = make_uint4(0, 0, 0, 0);
        if (act) e = queue[(qh + ei) & (QCAP - 1)];
        const uint32_t st = role ? e.y : e.x, P = role ? e.w : e.z;
#if defined(SALSA_ABL) && SALSA_ABL == 4          // (4: the queue is read, nothing else)
        npairs += e.x & 1u; qh += nb; qn -= nb; __syncwarp(); return;
#endif
        LineInfo L;
        L.id = -1; L.info = 0; L.noff = L.roff = 0; L.nlen = L.rlen = 0;
        long long sum = 0;
        const uint32_t p1 = P & 255u, p2 = (P >> 8) & 255u, p3 = (P >> 16) & 255u, rl = P >> 24;
        const uint32_t n1 = p2 - p1 - 1u, n2 = p3 - p2 - 1u;
        bool slow = act;
        if (act && P != 0u && st >= g_lo && st <= g_hi && n1 <= 12u && n2 <= 12u) {
            const unsigned long long g = (unsigned long long)st;      // offset from gsrc
            const uint32_t shn = (st & 3u) * 8u;
            const uint32_t *gp = gwords(gsrc, g);
            const uint32_t x0 = __ldg(gp), x1 = __ldg(gp + 1), x2 = __ldg(gp + 2), x3 = __ldg(gp + 3), x4 = __ldg(gp + 4);
            const uint4 mk = LM[p1 > 16u ? 16u : p1];
            const uint32_t w0 = __funnelshift_r(x0, x1, shn) & mk.x, w1 = __funnelshift_r(x1, x2, shn) & mk.y;
            const uint32_t w2 = __funnelshift_r(x2, x3, shn) & mk.z, w3 = __funnelshift_r(x3, x4, shn) & mk.w;
            // the first probe's latency hides behind the digit conversion
            const uint32_t hsh = salsa_name_hash(w0, w1, w2, w3, p1) & A.t.mask;
            const uint4 *slot = reinterpret_cast<const uint4 *>(&A.t.slots[hsh]);
            const uint4 k0 = __ldg(slot), m0 = __ldg(slot + 1);
            long long vs, ve;
            const bool dg1 = global_digits12(gsrc, LM, g + p2, n1, vs);
            const bool dg2 = global_digits12(gsrc, LM, g + p3, n2, ve);
            if (dg1 && dg2) {                        // else signs, junk: the full parser classifies it
                sum = vs + ve;
                if (p1 <= 16u && m0.x == p1 && k0.x == w0 && k0.y == w1 && k0.z == w2 && k0.w == w3) L.id = (int)m0.y;
                else if (p1 <= 16u && m0.x == 0xFFFFFFFFu) L.id = -1;
                else L.id = lookup_global(A.t, gsrc, w0, w1, w2, w3, g, p1, hsh);
                L.info = 4u;
                L.noff = lo + st; L.nlen = p1;
                L.roff = lo + st + p3 + 1u; L.rlen = rl;
                slow = false;
            }
        }
        if (slow) {                                  // results through local memory only on this path
            LineInfo T;
            long long tsum;
            parse_global(A, lo + st, T, tsum);
            L = T;
            sum = tsum;
        }
        if (L.id >= 0) {
            const ContigMeta m = A.t.meta[L.id];
            L.info |= ((m.flags & SALSA_CTG_HAS_LEN) ? 32u : 0u) | ((m.flags & SALSA_CTG_HAS_RE) ? 64u : 0u) |
                      (m.len_floor == 0 ? 128u : 0u) | (sum > m.len_floor ? 256u : 0u);
        }
#if defined(SALSA_ABL) && SALSA_ABL == 5          // (5: both lines parsed and looked up, nothing else)
        npairs += (uint32_t)(L.id + (int)sum + (int)L.info) & 1u; qh += nb; qn -= nb; __syncwarp(); return;
#endif
        // ---- read.split('/')[0] of both lines (make_links.py:78) ----
        const uint32_t oinfo = __shfl_xor_sync(0xffffffffu, L.info, 1);
        const int oid = __shfl_xor_sync(0xffffffffu, L.id, 1);
        const uint32_t orl = __shfl_xor_sync(0xffffffffu, L.rlen, 1);
        const bool pslow = ((L.info | oinfo) & 512u) != 0u;     // one of the pair's lines came through the full parser
        bool req;
        {
            // both lines plain: word by word, each lane its own line
            const unsigned long long r0 = (unsigned long long)st + p3 + 1u;
            const uint32_t nw = (act && !pslow) ? (rl + 3u) >> 2 : 0u;
            const uint32_t nwmax = __reduce_max_sync(0xffffffffu, nw);
            uint32_t diff = L.rlen ^ orl;
            const uint32_t *rp = gwords(gsrc, r0);
            const uint32_t rsh = ((uint32_t)r0 & 3u) * 8u;
#if defined(SALSA_ABL) && SALSA_ABL == 3          // (3: no read-name comparison)
            if (0)
#endif
            for (uint32_t j = 0; j < nwmax; j++) {
                uint32_t w = 0;
                if (j < nw) {
                    w = __funnelshift_r(__ldg(rp + j), __ldg(rp + j + 1), rsh);
                    const uint32_t left = rl - 4u * j;
                    if (left < 4u) w &= (1u << (8u * left)) - 1u;
                }
                diff |= w ^ __shfl_xor_sync(0xffffffffu, w, 1);
            }
            req = diff == 0u;
        }
        bool same_name = false;
        if (__any_sync(0xffffffffu, act && (pslow || (L.id < 0 && oid < 0)))) {
            // rare: compare byte by byte in global memory (a line from the full parser, or two unknown names)
            const unsigned long long oroff = __shfl_xor_sync(0xffffffffu, L.roff, 1), onoff = __shfl_xor_sync(0xffffffffu, L.noff, 1);
            const uint32_t onlen = __shfl_xor_sync(0xffffffffu, L.nlen, 1);
            if (act && !role) {
                if (pslow) req = L.rlen == orl && gbytes_equal(A, L.roff, oroff, orl);
                if (L.id < 0 && oid < 0) same_name = L.nlen == onlen && gbytes_equal(A, L.noff, onoff, onlen);
            }
        }
        // ---- the reference's body for one adjacent-line pair, in its order (make_links.py:70-144) ----
        if (act && !role) {
            const uint64_t line_off = file_base + e.x;
            const uint32_t cnf = L.info & 7u, pnf = oinfo & 7u, both = L.info | oinfo, all = L.info & oinfo;
            const bool same = (L.id >= 0 && oid >= 0) ? L.id == oid : ((L.id ^ oid) < 0 ? false : same_name);
            if (both & 16u) report(A.stats, line_off, SALSA_DATA_LINE_TOO_LONG);
            else if (pnf < 1u || cnf < 1u) report(A.stats, line_off, SALSA_DATA_INDEX_ERROR);                // :71
            else if (same) {}
            else if (pnf < 4u || cnf < 4u) report(A.stats, line_off, SALSA_DATA_INDEX_ERROR);                // :76-77
            else if (!req) {}                                                                                 // :78
            else if (both & 8u) report(A.stats, line_off, SALSA_DATA_VALUE_ERROR);                            // :81-82
            else if (L.id < 0 || oid < 0) report(A.stats, line_off, SALSA_DATA_KEY_ERROR);                    // :101-102
            else if (!(all & 32u)) report(A.stats, line_off, SALSA_DATA_KEY_ERROR);
            else if (both & 128u) report(A.stats, line_off, SALSA_DATA_ZERO_DIV);                             // :106-107
            else if (!(all & 64u)) report(A.stats, line_off, SALSA_DATA_KEY_ERROR);                           // :113
            else {
                // B <=> pos <= len/2 <=> start+end <= floor(len); E otherwise (:108-139); names ordered as strings (:90-96)
                const uint32_t ec = (L.info >> 8) & 1u, ep = (oinfo >> 8) & 1u;
                uint32_t first, second, ends;
                if (oid < L.id) { first = (uint32_t)oid; second = (uint32_t)L.id; ends = ep * 2u + ec; }
                else { first = (uint32_t)L.id; second = (uint32_t)oid; ends = ec * 2u + ep; }
                const uint64_t key = (((uint64_t)first * (uint64_t)A.t.n + second) << 2) | ends;
#if !defined(SALSA_ABL) || SALSA_ABL != 2         // (2: everything but the RED into the table)
                Emit::emit(EA, A, key, line_off);
#endif
                npairs++;
            }
        }
        qh += nb;
        qn -= nb;
        __syncwarp();
    };

    // ---- phase 1: nv (1..32) lines, entries lh .. lh + nv - 1 of the ring of line starts ----
    auto step = [&](uint32_t nv) {
        const bool actv = (uint32_t)lane < nv;
        const uint32_t start = lines[(lh + (uint32_t)lane) & (LCAP - 1)];
        Shp h;
        h.w0 = h.w1 = h.w2 = h.w3 = h.P = h.d = 0;
        if (actv) h = shape_of(s, spw, LM, start & (RING - 1), lim_r - start);
        // the line before: lane - 1's, and for lane 0 lane 31's of the step before (one rotation does both)
        const bool l31 = lane == 31;
        const int src = (lane + 31) & 31;
        Shp ph;
        ph.w0 = __shfl_sync(0xffffffffu, l31 ? hp.w0 : h.w0, src); ph.w1 = __shfl_sync(0xffffffffu, l31 ? hp.w1 : h.w1, src);
        ph.w2 = __shfl_sync(0xffffffffu, l31 ? hp.w2 : h.w2, src); ph.w3 = __shfl_sync(0xffffffffu, l31 ? hp.w3 : h.w3, src);
        ph.P = __shfl_sync(0xffffffffu, l31 ? hp.P : h.P, src); ph.d = __shfl_sync(0xffffffffu, l31 ? hp.d : h.d, src);
        const uint32_t pstart = __shfl_sync(0xffffffffu, l31 ? hp_start : start, src);
        const bool owned = actv && start > own_lo_r;
        bool cand = owned;
        if (!have_prev && lane == 0) {
            // a stream's first line start is known to be one only behind a newline of the halo
            if (owned) report(A.stats, file_base + start, SALSA_DATA_LINE_TOO_LONG);
            cand = false;
        }
        if (h.P != 0u && ph.P != 0u) {
            bool same = ((h.w0 ^ ph.w0) | (h.w1 ^ ph.w1) | (h.w2 ^ ph.w2) | (h.w3 ^ ph.w3)) == 0u;
            const uint32_t nl = h.P & 255u, pnl = ph.P & 255u;
            if (same && (nl | pnl) > 16u)        // a name longer than the window: the rest byte by byte (both lines are in the ring)
                same = nl == pnl && bytes_equal_back(s, (int)((pstart & (RING - 1)) + 16u), (int)((start & (RING - 1)) + 16u), (int)nl - 16);
            const bool rmay = ((h.P ^ ph.P) >> 24) == 0u && ((h.P >> 24) < 4u || h.d == ph.d);
            cand = cand && !same && rmay;              // names differ and the read names may be equal
        }
        const unsigned om = __ballot_sync(0xffffffffu, owned);
        const unsigned cm = __ballot_sync(0xffffffffu, cand);
        nlines += __popc(om);
        if (cand) queue[(qh + qn + __popc(cm & ltmask)) & (QCAP - 1)] = make_uint4(start, pstart, h.P, ph.P);
        qn += __popc(cm);
        hp = h;
        hp_start = start;
        if (nv != 32u) {               // a short step (end of the chunk, or the ring was full): its last line goes to lane 31
            const int last = (int)nv - 1;
            hp.w0 = __shfl_sync(0xffffffffu, h.w0, last); hp.w1 = __shfl_sync(0xffffffffu, h.w1, last);
            hp.w2 = __shfl_sync(0xffffffffu, h.w2, last); hp.w3 = __shfl_sync(0xffffffffu, h.w3, last);
            hp.P = __shfl_sync(0xffffffffu, h.P, last); hp.d = __shfl_sync(0xffffffffu, h.d, last);
            hp_start = __shfl_sync(0xffffffffu, start, last);
        }
        low = __shfl_sync(0xffffffffu, hp_start, 31);
        have_prev = true;
        lh += nv;
        __syncwarp();
    };

    // ---- classify 2 KiB from cpos on (64 bytes per lane): special bitmap + line starts ----
    auto classify = [&](auto cheap_tag) {
        constexpr bool CHEAP = decltype(cheap_tag)::value;
        const uint32_t c = cpos + 64u * (uint32_t)lane;
        const uint32_t ro = c & (RING - 1);
        const uint4 *src = reinterpret_cast<const uint4 *>(s + ro);
        const uint4 v0 = src[0], v1 = src[1], v2 = src[2], v3 = src[3];
        uint32_t s0, s1, s2, s3, n0, n1, n2, n3;
        masks16<CHEAP>(v0, s0, n0); masks16<CHEAP>(v1, s1, n1); masks16<CHEAP>(v2, s2, n2); masks16<CHEAP>(v3, s3, n3);
        uint32_t mlo = n0 | (n1 << 16), mhi = n2 | (n3 << 16);
        const uint2 sp = make_uint2(s0 | (s1 << 16), s2 | (s3 << 16));
        if (cpos >= nl_lo_r && cpos + (uint32_t)GRP <= nl_hi_r) {  // everywhere but in the chunk's first and last iterations
            *reinterpret_cast<uint2 *>(spw + (ro >> 5)) = sp;
            if (ro < (uint32_t)MIRROR) *reinterpret_cast<uint2 *>(spw + ((RING + ro) >> 5)) = sp;
        } else {
            if (c < wr_end_r) {                              // beyond the chunk the ring may hold older, live text
                *reinterpret_cast<uint2 *>(spw + (ro >> 5)) = sp;
                if (ro < (uint32_t)MIRROR) *reinterpret_cast<uint2 *>(spw + ((RING + ro) >> 5)) = sp;
            }
            // newlines outside [nl_lo, nl_hi) open no line of this chunk
            const long long a = (long long)nl_lo_r - (long long)c, b = (long long)nl_hi_r - (long long)c;
            const int a0 = a < 0 ? 0 : (a > 64 ? 64 : (int)a), b0 = b < 0 ? 0 : (b > 64 ? 64 : (int)b);
            const unsigned long long lowb = b0 >= 64 ? ~0ull : ((1ull << b0) - 1ull), lowa = a0 >= 64 ? ~0ull : ((1ull << a0) - 1ull);
            const unsigned long long keep = b0 > a0 ? (lowb & ~lowa) : 0ull;
            mlo &= (uint32_t)keep;
            mhi &= (uint32_t)(keep >> 32);
        }
        const unsigned hasl = __ballot_sync(0xffffffffu, mlo != 0u), hash = __ballot_sync(0xffffffffu, mhi != 0u);
        const unsigned multi = __ballot_sync(0xffffffffu, ((mlo & (mlo - 1u)) | (mhi & (mhi - 1u))) != 0u);
        if (multi == 0u) {                         // at most one newline per 32 bytes: rank by ballot
            const uint32_t rl = lt + __popc(hasl & ltmask) + __popc(hash & ltmask);
            if (mlo) lines[rl & (LCAP - 1)] = c + __ffs(mlo);
            if (mhi) lines[(rl + (mlo ? 1u : 0u)) & (LCAP - 1)] = c + 32u + __ffs(mhi);
            lt += __popc(hasl) + __popc(hash);
            cpos += (uint32_t)GRP;
        } else {
            // general case: inclusive scan of the counts; only the lanes whose newlines all fit the ring of line
            // starts are consumed, the next iteration starts at the first lane that was not
            const uint32_t room = (uint32_t)LCAP - (lt - lh);
            int cnt = __popc(mlo) + __popc(mhi), inc = cnt;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                int o = __shfl_up_sync(0xffffffffu, inc, d);
                if (lane >= d) inc += o;
            }
            const unsigned fit = __ballot_sync(0xffffffffu, (uint32_t)inc <= room);
            const int ncons = fit == 0xFFFFFFFFu ? 32 : __ffs(~fit) - 1;       // a prefix of the lanes (inc is monotone), >= 1
            uint32_t pos = lt + (uint32_t)(inc - cnt);
            if (lane < ncons) {
                while (mlo) {
                    const int b = __ffs(mlo);
                    mlo &= mlo - 1u;
                    lines[pos & (LCAP - 1)] = c + (uint32_t)b;
                    pos++;
                }
                while (mhi) {
                    const int b = __ffs(mhi);
                    mhi &= mhi - 1u;
                    lines[pos & (LCAP - 1)] = c + 32u + (uint32_t)b;
                    pos++;
                }
            }
            lt += (uint32_t)__shfl_sync(0xffffffffu, inc, ncons - 1);
            cpos += 64u * (uint32_t)ncons;
        }
        __syncwarp();
    };

    // ---- the chunk ----
    for (;;) {
        // classify until 32 lines can be looked at, or until nothing more can be classified right now
        uint32_t cnt;
        bool canstep;
        for (;;) {
            cnt = lt - lh;
            canstep = cnt >= 32u && cpos - lines[(lh + 31u) & (LCAP - 1)] >= over_eff;
            if (canstep || cpos >= rel_end_r || cnt > (uint32_t)LCAP - 64u) break;
            issue_allowed();                                  // usually one group: what the last iteration or step made room for
            const uint32_t want = cpos + (uint32_t)GRP < rel_end_r ? cpos + (uint32_t)GRP : rel_end_r;
            const uint32_t g = (want - 1u) / GRP;             // the group this iteration reads into
            if (g >= gi) break;                               // the ring is full
            if (g >= gw) wait_for(g);
            if (cheap) classify(CheapTest());
            else classify(ExactTest());
            if (cpos >= rel_end_r) over_eff = 0u;
        }
        uint32_t nv = 32u;
        if (!canstep) {
            // ---- the rare paths: the end of the chunk, a line longer than the ring ----
            if (cpos < rel_end_r) {
                // no room in the ring, because of one very long line: look at what can be looked at, then let go of the line before it
                const uint32_t vi = lines[(lh + (uint32_t)lane) & (LCAP - 1)];
                nv = __popc(__ballot_sync(0xffffffffu, (uint32_t)lane < cnt && cpos - vi >= (uint32_t)OVER));
                if (!nv) {
                    const uint32_t st0 = cnt ? lines[lh & (LCAP - 1)] : cpos;      // nothing older than this is needed any more
                    if (low != st0 || hp.P) { hp.P = 0u; low = st0; continue; }
                    if (lane == 0) report(A.stats, file_base + cpos, SALSA_DATA_LINE_TOO_LONG);     // cannot happen: RING >> OVER + 2 * GRP
                    break;
                }
            } else if (cnt) nv = cnt < 32u ? cnt : 32u;     // the chunk's last lines
            else break;
        }
        step(nv);
        while (qn >= (uint32_t)QPAIRS) flush(QPAIRS);
    }
    while (qn) flush(qn < (uint32_t)QPAIRS ? qn : (uint32_t)QPAIRS);
    __syncwarp();
    }   // chunks

    // ---- statistics: one atomic pair per warp ----
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) npairs += __shfl_xor_sync(0xffffffffu, npairs, d);
    if (lane == 0) {
        if (nlines) atomicAdd((unsigned long long *)&A.stats[1], (unsigned long long)nlines);
        if (npairs) atomicAdd((unsigned long long *)&A.stats[2], (unsigned long long)npairs);
    }
}

template <class Emit>
int launch_stream(const typename Emit::Args &EA, const DevTable &t, const void *d_bed, size_t nbytes, uint64_t base_offset,
                  uint64_t *stats, int scan_mode, int names_punct, int *d_mode, cudaStream_t st)
{
    if (nbytes == 0) return SALSA_OK;
    ScanArgs A;
    uintptr_t p = (uintptr_t)d_bed;
    A.mis = (uint32_t)(p & 15u);
    A.abase = (const uint8_t *)(p - A.mis);
    A.dend = (uint64_t)A.mis + nbytes;
    A.base_offset = base_offset;
    A.t = t;
    A.stats = stats;
    static int ctas_per_dev[64] = {0};
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) SALSA_FAIL(SALSA_ERR_ARG, "device index %d out of range", dev);
    if (!ctas_per_dev[dev]) {
        int sms = 0, per = 0;
        CUDA_TRY(cudaFuncSetAttribute(bed_stream_kernel<Emit>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CTA_SMEM));
        CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, bed_stream_kernel<Emit>, WARPS * 32, CTA_SMEM));
        if (per < 1) SALSA_FAIL(SALSA_ERR_CUDA, "bed_stream_kernel does not fit an SM (%zu bytes of shared memory)", CTA_SMEM);
        ctas_per_dev[dev] = sms * per;
    }
    CUDA_TRY(cudaFuncSetAttribute(bed_stream_kernel<Emit>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CTA_SMEM));
    const uint64_t nwarps = (uint64_t)ctas_per_dev[dev] * WARPS;
    uint64_t cb = (A.dend + nwarps - 1) / nwarps;       // every resident warp gets at least one chunk ...
    if (cb < CHUNK_MIN) cb = CHUNK_MIN;
    if (cb > CHUNK_MAX) cb = CHUNK_MAX;                 // ... and large inputs are dealt out in chunks of CHUNK_MAX
    cb = (cb + 1023) & ~1023ull;                        // chunks start on 1 KiB: ring offsets, stages and iterations stay aligned
    const uint64_t nchunks = (A.dend + cb - 1) / cb;
    if (nchunks > 0x7fffffffull) SALSA_FAIL(SALSA_ERR_ARG, "BED chunk too large (%zu bytes)", nbytes);
    uint64_t ctas = (nchunks + WARPS - 1) / WARPS;
    if (ctas > (uint64_t)ctas_per_dev[dev]) ctas = (uint64_t)ctas_per_dev[dev];
    // large inputs: the counter hands out quarter chunks, four at a time until the text is nearly used up, then singly
    uint64_t unit = cb, take_big = 1, tail_units = 0;
    if (cb == CHUNK_MAX && nchunks > 8 * nwarps) {
        unit = cb / 4;
        take_big = 4;
        tail_units = 8 * nwarps;                        // the last two chunks' worth per warp
    }
    const uint64_t nunits = (A.dend + unit - 1) / unit;
    if (nunits > 0x3fffffffull) SALSA_FAIL(SALSA_ERR_ARG, "BED chunk too large (%zu bytes)", nbytes);
    // scan_mode: 0 = the sample kernel decides, 1 = exact special-byte test, 2 = cheap one
    if (scan_mode == 0) bed_sample_kernel<<<1, 256, 0, st>>>(A, names_punct, d_mode);
    else CUDA_TRY(cudaMemsetAsync(d_mode, scan_mode == 2 ? 1 : 0, 1, st));
    CUDA_TRY(cudaMemsetAsync(d_mode + 1, 0, sizeof(int), st));          // the chunk counter
    bed_stream_kernel<Emit><<<(unsigned)ctas, WARPS * 32, CTA_SMEM, st>>>(EA, A, (uint32_t)unit, (uint32_t)nunits, (uint32_t)take_big, (uint32_t)tail_units,
                                                                         d_mode);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}

}   // namespace bedstream
