// bed_scan.cuh -- the shared BED text scanner: tiles, line starts, per-line records,
// adjacent-line pairing.  Instantiated with a Policy by bed_ingest.cu (link
// histogram, make_links.py:60-146) and coverage.cu (physical-coverage diff arrays,
// break_contigs_start.cpp:307-385 / break_contigs.cpp:199-254).
//
// Shape of the kernel:
//   * the BED is cut into TILE-byte tiles; one CTA stages its tile plus a
//     SALSA_MAX_LINE halo on either side in shared memory with a single 1-D
//     TMA bulk copy (cp.async.bulk ... mbarrier::complete_tx);
//   * a tile OWNS the lines whose first byte lies in (t0, t1]; it evaluates the
//     adjacent-line pairs whose second line it owns, so every pair is
//     evaluated exactly once and the line at offset 0 only ever plays "previous";
//   * CLASSIFY (cooperative, 32 bytes per lane and iteration, exact SWAR byte
//     tests + dp4a bit packing): a bitmap of "special" bytes (<= 0x20 or '/') for
//     the whole staged region, and the newline positions of the owned region,
//     compacted into per-warp lists of line starts (ballot rank; a general warp
//     scan when 32 bytes hold two newlines);
//   * PARSE (one thread per line, branch-free fast path): the first four set
//     bits of the special bitmap are the field boundaries; the contig name is
//     looked up from a 16-byte unaligned window, start/end are converted 12
//     digits at a time (xor 0x30, SWAR digit check, dp4a), the read name is
//     packed into 8 words.  Anything the fast path cannot prove well-formed
//     (odd whitespace, signs, long fields, '/' in a contig name ...) goes
//     through the byte-loop parser parse_line(), which implements str.split()
//     in full -- results are identical, only speed differs;
//   * the record travels to the next lane by warp shuffle -- the "previous
//     line" -- and Policy::pair() decides what the adjacent pair contributes;
//   * policies with LAZY = true (the link histogram) run that in three phases:
//     every line only gets its field layout (line_shape); pairs that pair()
//     provably neither counts nor reports are dropped there, the rest is queued
//     and gets the full parse in batches of 32 per warp, then across the CTA
//     (process_range, "Three phases").
#pragma once
#include "common.cuh"

namespace bedscan {

#ifndef SALSA_TILE
#define SALSA_TILE 32768
#endif
#ifndef SALSA_MINBLOCKS
#define SALSA_MINBLOCKS 5
#endif
constexpr int TILE = SALSA_TILE;
constexpr int HALO = SALSA_MAX_LINE;   // 1024
constexpr int NTHREADS = 256;
constexpr int NWARPS = NTHREADS / 32;
constexpr int LCAP = 2048;             // line starts per scan range
constexpr int WCAP = LCAP / NWARPS;    // per-warp segment
constexpr int SUBTILE = 2048;          // fallback range: <= 2048 line starts by construction
constexpr int SPARE_OFF = HALO + TILE + HALO + 32;      // one spare line (avoid mode: a line fetched from before the tile)
constexpr int SMEM_BYTES = SPARE_OFF + HALO + 64;
constexpr int SPWORDS = SPARE_OFF / 32 + 8;   // special-byte bitmap, 1 bit per staged byte (+ guard words)
constexpr int RW = 8;                  // read-name words carried per record (32 bytes)
constexpr int QCAP = 72;               // lazy warp loop: per-warp queue of pending line indices (< 32 left + <= 31 new) + 8 bytes of bookkeeping

struct ScanArgs {
    const uint8_t *abase;     // 16-byte aligned base (<= caller pointer)
    uint32_t mis;             // caller pointer - abase  (0..15)
    uint64_t dend;            // mis + nbytes : one past the last valid byte (abase coordinates)
    uint64_t base_offset;     // file offset of the caller's byte 0 (error reports)
    DevTable t;
    uint64_t *stats;
};

struct Rec {
    int id;               // contig id, -1 unknown
    int s_lo, s_hi;       // start (int64)
    int e_lo, e_hi;       // end (int64)
    uint32_t name;        // noff | nlen << 16   (smem offsets)
    uint32_t read;        // roff | rcmp << 16: rcmp = number of read-name bytes the policy compares
    uint32_t flags;       // nf (bits 0-2) | numbad (bit 3) | toolong (bit 4) | whole read token length << 8 (NEED_RFULL policies)
    uint32_t rw[RW];      // first min(rcmp, 32) read-name bytes, zero padded
};

// scratch shared by one CTA
struct Smem {
    uint8_t *s;           // staged bytes
    uint32_t *sp;         // special bitmap
    uint16_t *seg;        // per-warp newline segments
    uint16_t *lines;      // compacted line starts
    uint32_t (*lowmask)[RW];   // lowmask[e][j]: bytes of word j that belong to the low e bytes of a window
    Rec *carry;
    int *wcount;
    int *flagp;
    uint8_t *queue;       // lazy policies: per-warp ring (QCAP bytes) of line indices waiting for the full parse
    // avoid mode only (Policy::AVOID)
    int *idv;             // per line entry: contig id | IDV_INVOLVED, -1 unknown
    Rec *ext;             // record of the effective previous line when it lies before the tile
    int *extinfo;         // [0] entry 0 was skipped, [1] contig id of *ext
};

constexpr int IDV_INVOLVED = 0x40000000;
constexpr int IDV_KEPT = 0x20000000;       // the line is not skipped by the avoid rule (current iterate of the fixed point below)
constexpr int IDV_NEXT = 0x10000000;       // ... its next iterate
__device__ __forceinline__ int idv_id(int v) { return v < 0 ? -1 : (v & 0x0FFFFFFF); }
__device__ __forceinline__ bool idv_kept(int v) { return v < 0 || (v & IDV_KEPT); }
__device__ __forceinline__ bool idv_inv(int v) { return v >= 0 && (v & IDV_INVOLVED); }

__device__ __forceinline__ long long rec_start(const Rec &r)
{
    return (long long)(((unsigned long long)(uint32_t)r.s_hi << 32) | (uint32_t)r.s_lo);
}
__device__ __forceinline__ long long rec_end(const Rec &r)
{
    return (long long)(((unsigned long long)(uint32_t)r.e_hi << 32) | (uint32_t)r.e_lo);
}

__device__ __forceinline__ bool is_ws(uint32_t c) { return c == 32u || (c - 9u) <= 4u; }

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- exact SWAR byte tests: 0x80 in every byte that satisfies the test ----
__device__ __forceinline__ uint32_t eq_bytes(uint32_t w, uint32_t c4)
{
    uint32_t x = w ^ c4;
    uint32_t t = (x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu;
    return ~(t | x) & 0x80808080u;
}
__device__ __forceinline__ uint32_t le_bytes_0x20(uint32_t w)
{
    uint32_t t = (w & 0x7F7F7F7Fu) + 0x5F5F5F5Fu;     // bit 7 set iff (b & 0x7f) > 0x20
    return ~(t | w) & 0x80808080u;
}
// four 0x80-flag words -> 16-bit mask, bit k = byte k of the 16-byte chunk
__device__ __forceinline__ uint32_t pack16(uint32_t z0, uint32_t z1, uint32_t z2, uint32_t z3)
{
    uint32_t lo = __dp4a(z0, 0x08040201u, 0u);
    lo = __dp4a(z1, 0x80402010u, lo);                 // 128 * (nib0 | nib1 << 4)
    uint32_t hi = __dp4a(z2, 0x08040201u, 0u);
    hi = __dp4a(z3, 0x80402010u, hi);
    return (lo >> 7) | (hi << 1);
}
__device__ __forceinline__ uint32_t nl16_of(const uint4 &v)
{
    return pack16(eq_bytes(v.x, 0x0A0A0A0Au), eq_bytes(v.y, 0x0A0A0A0Au), eq_bytes(v.z, 0x0A0A0A0Au), eq_bytes(v.w, 0x0A0A0A0Au));
}
__device__ __forceinline__ uint32_t sp16_of(const uint4 &v)
{
    return pack16(le_bytes_0x20(v.x) | eq_bytes(v.x, 0x2F2F2F2Fu), le_bytes_0x20(v.y) | eq_bytes(v.y, 0x2F2F2F2Fu),
                  le_bytes_0x20(v.z) | eq_bytes(v.z, 0x2F2F2F2Fu), le_bytes_0x20(v.w) | eq_bytes(v.w, 0x2F2F2F2Fu));
}

__device__ __forceinline__ void report(uint64_t *stats, uint64_t file_off, int code)
{
    atomicMin((unsigned long long *)&stats[0], (unsigned long long)((file_off << 8) | (uint64_t)code));
}

// ---- contig name -> id ------------------------------------------------
__device__ __forceinline__ int lookup_words(const DevTable &t, const uint8_t *s, uint32_t w0, uint32_t w1, uint32_t w2,
                                            uint32_t w3, int noff, int nlen)
{
    uint32_t h = salsa_name_hash(w0, w1, w2, w3, (uint32_t)nlen) & t.mask;
    for (;;) {
        const uint4 *sp = reinterpret_cast<const uint4 *>(&t.slots[h]);
        uint4 k = __ldg(sp);
        uint4 m = __ldg(sp + 1);     // len, id, blob_off, pad
        if (m.x == 0xFFFFFFFFu) return -1;
        if (m.x == (uint32_t)nlen && k.x == w0 && k.y == w1 && k.z == w2 && k.w == w3) {
            if (nlen <= 16) return (int)m.y;
            const char *b = t.blob + m.z;
            bool eq = true;
            for (int i = 16; i < nlen; i++)
                if ((uint8_t)__ldg(&b[i]) != s[noff + i]) { eq = false; break; }
            if (eq) return (int)m.y;
        }
        h = (h + 1) & t.mask;
    }
}

// 16 bytes at smem offset `off` (any alignment) -> 4 words, bytes >= n cleared
__device__ __forceinline__ void window4(const Smem &S, int off, int n, uint32_t &w0, uint32_t &w1, uint32_t &w2, uint32_t &w3)
{
    const uint32_t *s32 = reinterpret_cast<const uint32_t *>(S.s);
    int a = off >> 2, sh = (off & 3) * 8;
    uint32_t x0 = s32[a], x1 = s32[a + 1], x2 = s32[a + 2], x3 = s32[a + 3], x4 = s32[a + 4];
    const uint32_t *mk = S.lowmask[n > 16 ? 16 : n];
    w0 = __funnelshift_r(x0, x1, sh) & mk[0];
    w1 = __funnelshift_r(x1, x2, sh) & mk[1];
    w2 = __funnelshift_r(x2, x3, sh) & mk[2];
    w3 = __funnelshift_r(x3, x4, sh) & mk[3];
}

__device__ __forceinline__ int lookup_name(const DevTable &t, const Smem &S, int noff, int nlen)
{
    uint32_t w0, w1, w2, w3;
    window4(S, noff, nlen, w0, w1, w2, w3);
    return lookup_words(t, S.s, w0, w1, w2, w3, noff, nlen);
}

// read-name bytes [roff, roff + min(n, 32)) -> rw[], zero padded
__device__ __forceinline__ void pack_read(const Smem &S, int roff, int n, uint32_t *rw)
{
    const uint32_t *s32 = reinterpret_cast<const uint32_t *>(S.s);
    int a = roff >> 2, sh = (roff & 3) * 8;
    const uint32_t *mk = S.lowmask[n > 32 ? 32 : n];
    uint32_t x = s32[a];
#pragma unroll
    for (int j = 0; j < RW; j++) {
        uint32_t y = s32[a + j + 1];
        rw[j] = __funnelshift_r(x, y, sh) & mk[j];
        x = y;
    }
}

// the n (1..12) decimal digits that END just before smem offset b -> value; false if a byte is not a digit
__device__ __forceinline__ bool digits12(const Smem &S, int b, int n, long long &val)
{
    const uint32_t *s32 = reinterpret_cast<const uint32_t *>(S.s);
    int q = b - 12;
    int a = q >> 2, sh = (q & 3) * 8;
    uint32_t x0 = s32[a], x1 = s32[a + 1], x2 = s32[a + 2], x3 = s32[a + 3];
    const uint32_t *mk = S.lowmask[12 - n];           // the low 12 - n bytes of the window precede the field
    uint32_t u0 = (__funnelshift_r(x0, x1, sh) ^ 0x30303030u) & ~mk[0];
    uint32_t u1 = (__funnelshift_r(x1, x2, sh) ^ 0x30303030u) & ~mk[1];
    uint32_t u2 = (__funnelshift_r(x2, x3, sh) ^ 0x30303030u) & ~mk[2];
    // byte ^ 0x30 <= 9 exactly for '0'..'9'
    uint32_t bad = ((((u0 & 0x7F7F7F7Fu) + 0x76767676u) | u0) | (((u1 & 0x7F7F7F7Fu) + 0x76767676u) | u1) |
                    (((u2 & 0x7F7F7F7Fu) + 0x76767676u) | u2)) & 0x80808080u;
    uint32_t v0 = __dp4a(u0, 0x0000010Au, 0u) * 100u + __dp4a(u0, 0x010A0000u, 0u);
    uint32_t v1 = __dp4a(u1, 0x0000010Au, 0u) * 100u + __dp4a(u1, 0x010A0000u, 0u);
    uint32_t v2 = __dp4a(u2, 0x0000010Au, 0u) * 100u + __dp4a(u2, 0x010A0000u, 0u);
    val = (long long)((unsigned long long)(v0 * 10000u + v1) * 10000ull + v2);
    return bad == 0;
}

// ---- one line -> record, byte-loop parser: str.split() in full (make_links.py:62) ----
// p: smem index of the line's first byte; sent: smem index of the '\n' sentinel that ends
// the staged bytes; trunc: the sentinel stands for a halo cut, not for the end of the
// data.  Numbers are strict [+-]?[0-9]{1,18}.  *rfull = length of the whole read token.
static __device__ __noinline__ Rec parse_line(const DevTable &t, const Smem &S, int p, int sent, bool trunc, int *rslash, int *rfull)
{
    const uint8_t *s = S.s;
    Rec r;
    r.id = -1; r.s_lo = r.s_hi = r.e_lo = r.e_hi = 0; r.name = 0; r.read = 0; r.flags = 0;
#pragma unroll
    for (int j = 0; j < RW; j++) r.rw[j] = 0;
    *rslash = 0; *rfull = 0;
    uint32_t c = s[p];
    while (c != '\n' && is_ws(c)) c = s[++p];
    int nf = 0;
    uint32_t bad = 0;
    if (c != '\n') {
        int noff = p;
        do { c = s[++p]; } while (!is_ws(c));
        int nlen = p - noff;
        r.name = (uint32_t)noff | ((uint32_t)nlen << 16);
        r.id = lookup_name(t, S, noff, nlen);
        nf = 1;
        long long val[2] = {0, 0};
#pragma unroll
        for (int f = 0; f < 2; f++) {
            if (nf == 1 + f) {
                while (c != '\n' && is_ws(c)) c = s[++p];
                if (c != '\n') {
                    nf++;
                    bool neg = (c == '-');
                    if (c == '-' || c == '+') c = s[++p];
                    unsigned long long v = 0;
                    int nd = 0;
                    uint32_t d = c - '0';
                    while (d <= 9u) { v = v * 10ull + d; nd++; c = s[++p]; d = c - '0'; }
                    if (nd == 0 || nd > 18 || !is_ws(c)) bad = 8u;
                    while (!is_ws(c)) c = s[++p];
                    val[f] = neg ? -(long long)v : (long long)v;
                }
            }
        }
        r.s_lo = (int)(uint32_t)(unsigned long long)val[0];
        r.s_hi = (int)(uint32_t)((unsigned long long)val[0] >> 32);
        r.e_lo = (int)(uint32_t)(unsigned long long)val[1];
        r.e_hi = (int)(uint32_t)((unsigned long long)val[1] >> 32);
        if (nf == 3) {
            while (c != '\n' && is_ws(c)) c = s[++p];
            if (c != '\n') {
                nf = 4;
                int roff = p;
                while (c != '/' && !is_ws(c)) c = s[++p];      // read.split('/')[0]
                *rslash = p - roff;
                while (!is_ws(c)) c = s[++p];                  // rest of the token
                *rfull = p - roff;
                r.read = (uint32_t)roff;
            }
        }
    }
    r.flags = (uint32_t)nf | bad | ((trunc && p >= sent) ? 16u : 0u);
    return r;
}

// 64 bits of the special bitmap starting at smem index p
__device__ __forceinline__ unsigned long long sp_window(const Smem &S, int p)
{
    int i = p >> 5, sh = p & 31;
    uint32_t a0 = S.sp[i], a1 = S.sp[i + 1], a2 = S.sp[i + 2];
    return (unsigned long long)__funnelshift_r(a0, a1, sh) | ((unsigned long long)__funnelshift_r(a1, a2, sh) << 32);
}

// The first four special bytes of the line at p (offsets relative to p).  A name longer than 16 bytes gets a
// second window centred on its separator, so names up to 63 bytes and 63 more bytes of fields are in reach.
// On return M's lowest set bit is p4's (relative to `base`): callers continue with M &= M - 1.
__device__ __forceinline__ bool bounds4(const Smem &S, int p, int &p1, int &p2, int &p3, int &p4, unsigned long long &M, int &base)
{
    const unsigned long long M0 = sp_window(S, p);
    M = M0;
    base = 0;
    if (M & 1ull) return false;                          // leading whitespace / empty line
    // common case first, as straight-line code: short name, four specials inside the one window
    p1 = __ffsll((long long)M) - 1;
    M &= M - 1;
    p2 = __ffsll((long long)M) - 1;
    M &= M - 1;
    p3 = __ffsll((long long)M) - 1;
    M &= M - 1;
    if (!(M != 0ull && p1 <= 16)) {
        if (!M0) return false;
        M = M0;
        if (p1 > 16) {                                   // long name: a window centred on its separator
            base = p1;
            M = sp_window(S, p + p1);
        }
        M &= M - 1;
        p2 = base + __ffsll((long long)M) - 1;
        M &= M - 1;
        if (!M) return false;
        p3 = base + __ffsll((long long)M) - 1;
        M &= M - 1;
        if (!M) {                                        // long read name: a window centred on its separator
            base = p3;
            M = sp_window(S, p + p3);
            M &= M - 1;
            if (!M) return false;
        }
    }
    p4 = base + __ffsll((long long)M) - 1;
    return true;
}

// ---- one line -> record, branch-free fast path.  Returns false when the line is not
// provably of the plain form  name SEP digits SEP digits SEP read[/...]  (SEP = one tab
// or space) with name <= 63, digits <= 12 bytes and all of it inside the staged bytes.
template <bool NEED_RFULL>
__device__ __forceinline__ bool parse_fast(const DevTable &t, const Smem &S, int p, int sent, Rec &r, int &rslash, int &rfull)
{
    const uint8_t *s = S.s;
    unsigned long long M;
    int p1, p2, p3, p4, base;
    if (!bounds4(S, p, p1, p2, p3, p4, M, base)) return false;   // leading whitespace, empty line, < 4 specials in reach
    uint32_t c1 = s[p + p1], c2 = s[p + p2], c3 = s[p + p3], c4 = s[p + p4];
    bool ok = (c1 == 9u || c1 == 32u) && (c2 == 9u || c2 == 32u) && (c3 == 9u || c3 == 32u) &&
              (c4 == 47u || c4 == 9u || c4 == 32u || c4 == 10u);
    int n1 = p2 - p1 - 1, n2 = p3 - p2 - 1, rl = p4 - p3 - 1;
    ok = ok && n1 >= 1 && n1 <= 12 && n2 >= 1 && n2 <= 12 && rl >= 1 && (p + p4) < sent;
    rslash = rl;
    rfull = rl;
    if (NEED_RFULL) {
        if (c4 == 47u) {                                 // token continues after the first '/'
            M &= M - 1;
            if (!M) return false;
            int p5 = base + __ffsll((long long)M) - 1;
            uint32_t c5 = s[p + p5];
            ok = ok && (c5 == 9u || c5 == 32u || c5 == 10u) && (p + p5) < sent;
            rfull = p5 - p3 - 1;
        }
    }
    if (!ok) return false;
    // issue the first name-table probe now; its latency hides behind the digit conversion
    uint32_t w0, w1, w2, w3;
    window4(S, p, p1, w0, w1, w2, w3);
    uint32_t h = salsa_name_hash(w0, w1, w2, w3, (uint32_t)p1) & t.mask;
    const uint4 *slot = reinterpret_cast<const uint4 *>(&t.slots[h]);
    uint4 k0 = __ldg(slot), m0 = __ldg(slot + 1);
    long long vs, ve;
    bool d1 = digits12(S, p + p2, n1, vs);
    bool d2 = digits12(S, p + p3, n2, ve);
    if (!(d1 && d2)) return false;                       // signs, junk: let the full parser classify it
    r.name = (uint32_t)p | ((uint32_t)p1 << 16);
    if (p1 <= 16 && m0.x == (uint32_t)p1 && k0.x == w0 && k0.y == w1 && k0.z == w2 && k0.w == w3) r.id = (int)m0.y;   // exact
    else if (p1 > 16) r.id = lookup_words(t, s, w0, w1, w2, w3, p, p1);                                   // long name: verified against the blob
    else if (m0.x == 0xFFFFFFFFu) r.id = -1;
    else r.id = lookup_words(t, s, w0, w1, w2, w3, p, p1);                                                // collision: probe on
    r.s_lo = (int)(uint32_t)(unsigned long long)vs; r.s_hi = (int)(uint32_t)((unsigned long long)vs >> 32);
    r.e_lo = (int)(uint32_t)(unsigned long long)ve; r.e_hi = (int)(uint32_t)((unsigned long long)ve >> 32);
    r.read = (uint32_t)(p + p3 + 1);
    r.flags = 4u;
    return true;
}

__device__ __forceinline__ bool bytes_equal_back(const uint8_t *s, int a, int b, int n)
{
    for (int i = n - 1; i >= 0; i--)
        if (s[a + i] != s[b + i]) return false;
    return true;
}

// read-name equality over the policy's compare length (first 32 bytes from the packed words)
__device__ __forceinline__ bool reads_equal(const uint8_t *s, const Rec &pv, const Rec &cu)
{
    uint32_t pl = pv.read >> 16, cl = cu.read >> 16;
    uint32_t d = pl ^ cl;
#pragma unroll
    for (int j = 0; j < RW; j++) d |= pv.rw[j] ^ cu.rw[j];
    if (d) return false;
    if (cl <= 32u) return true;
    int pa = (int)(pv.read & 0xFFFF) + 32, ca = (int)(cu.read & 0xFFFF) + 32;
    return bytes_equal_back(s, pa, ca, (int)cl - 32);
}

// contig equality of two adjacent records by NAME (ids when both are known)
__device__ __forceinline__ bool same_contig(const uint8_t *s, const Rec &pv, const Rec &cu)
{
    if (pv.id >= 0 && cu.id >= 0) return pv.id == cu.id;
    if ((pv.id ^ cu.id) < 0) return false;          // one known, one unknown
    int pl = pv.name >> 16, cl = cu.name >> 16;
    return pl == cl && bytes_equal_back(s, pv.name & 0xFFFF, cu.name & 0xFFFF, cl);
}

__device__ __forceinline__ Rec rec_shfl_idx(const Rec &r, int src)
{
    Rec o;
    o.id = __shfl_sync(0xffffffffu, r.id, src);
    o.s_lo = __shfl_sync(0xffffffffu, r.s_lo, src);
    o.s_hi = __shfl_sync(0xffffffffu, r.s_hi, src);
    o.e_lo = __shfl_sync(0xffffffffu, r.e_lo, src);
    o.e_hi = __shfl_sync(0xffffffffu, r.e_hi, src);
    o.name = __shfl_sync(0xffffffffu, r.name, src);
    o.read = __shfl_sync(0xffffffffu, r.read, src);
    o.flags = __shfl_sync(0xffffffffu, r.flags, src);
#pragma unroll
    for (int j = 0; j < RW; j++) o.rw[j] = __shfl_sync(0xffffffffu, r.rw[j], src);
    return o;
}

__device__ __forceinline__ Rec rec_shfl_up(const Rec &r)
{
    Rec o;
    o.id = __shfl_up_sync(0xffffffffu, r.id, 1);
    o.s_lo = __shfl_up_sync(0xffffffffu, r.s_lo, 1);
    o.s_hi = __shfl_up_sync(0xffffffffu, r.s_hi, 1);
    o.e_lo = __shfl_up_sync(0xffffffffu, r.e_lo, 1);
    o.e_hi = __shfl_up_sync(0xffffffffu, r.e_hi, 1);
    o.name = __shfl_up_sync(0xffffffffu, r.name, 1);
    o.read = __shfl_up_sync(0xffffffffu, r.read, 1);
    o.flags = __shfl_up_sync(0xffffffffu, r.flags, 1);
#pragma unroll
    for (int j = 0; j < RW; j++) o.rw[j] = __shfl_up_sync(0xffffffffu, r.rw[j], 1);
    return o;
}

// one line -> record: fast path, else the full parser; read name packed per policy
template <class Policy>
__device__ __forceinline__ Rec parse_any(const ScanArgs &A, const Smem &S, int lstart, int sent, bool trunc, bool allow_fast)
{
    Rec rec;
    rec.id = -1; rec.s_lo = rec.s_hi = rec.e_lo = rec.e_hi = 0; rec.name = rec.read = 0; rec.flags = 0;
#pragma unroll
    for (int j = 0; j < RW; j++) rec.rw[j] = 0;
    int rslash = 0, rfull = 0;
    if (!allow_fast || !parse_fast<Policy::NEED_RFULL>(A.t, S, lstart, sent, rec, rslash, rfull))
        rec = parse_line(A.t, S, lstart, sent, trunc, &rslash, &rfull);
    if ((rec.flags & 7u) == 4u) {
        int rcmp = Policy::read_cmp_len(rslash, rfull);
        int roff = (int)(rec.read & 0xFFFF);
        rec.read = (uint32_t)roff | ((uint32_t)rcmp << 16);
        if (Policy::NEED_RFULL) rec.flags |= (uint32_t)rfull << 8;
        pack_read(S, roff, rcmp, rec.rw);
    }
    return rec;
}

// ---- lazy warp loop, phase 1: a line's field layout without converting anything ----
// nl == NOT_PLAIN: the line is not provably of the plain form (the conditions of parse_fast minus the digit
// checks; names up to what fits the 64-byte window).  Otherwise: name
// length, first 16 name bytes (zero padded), read-prefix length and the last min(8, length) bytes of the read prefix.
// Equal names <=> equal (nl, bytes); equal read prefixes => equal (rl, d0, d1).
struct Shape {
    uint32_t w0, w1, w2, w3, nl, rl, d0, d1;
};
constexpr uint32_t NOT_PLAIN = 0xFFFFu;

__device__ __forceinline__ Shape line_shape(const Smem &S, int p, int sent)
{
    Shape h;
    h.w0 = h.w1 = h.w2 = h.w3 = 0; h.nl = NOT_PLAIN; h.rl = h.d0 = h.d1 = 0;
    const uint8_t *s = S.s;
    unsigned long long M;
    int p1, p2, p3, p4, base;
    if (!bounds4(S, p, p1, p2, p3, p4, M, base)) return h;
    uint32_t c1 = s[p + p1], c2 = s[p + p2], c3 = s[p + p3], c4 = s[p + p4];
    int rl = p4 - p3 - 1;
    bool ok = (c1 == 9u || c1 == 32u) && (c2 == 9u || c2 == 32u) && (c3 == 9u || c3 == 32u) &&
              (c4 == 47u || c4 == 9u || c4 == 32u || c4 == 10u) && p2 - p1 > 1 && p3 - p2 > 1 && rl >= 1 && (p + p4) >= 8 && (p + p4) < sent;
    if (!ok) return h;
    window4(S, p, p1, h.w0, h.w1, h.w2, h.w3);
    const uint32_t *s32 = reinterpret_cast<const uint32_t *>(s);
    int q = p + p4 - 8, a = q >> 2, shq = (q & 3) * 8;
    uint32_t x0 = s32[a], x1 = s32[a + 1], x2 = s32[a + 2];
    h.d0 = __funnelshift_r(x0, x1, shq);
    h.d1 = __funnelshift_r(x1, x2, shq);
    if (rl < 8) {                                        // short read names: only their own bytes count
        const uint32_t *mk = S.lowmask[8 - rl];
        h.d0 &= ~mk[0];
        h.d1 &= ~mk[1];
    }
    h.nl = (uint32_t)p1;
    h.rl = (uint32_t)rl;
    return h;
}

// ---- avoid list (links_avoid_iteration_N, make_links.py:29-35,72-74) ------
// `continue` on an avoided contig pair leaves prev_line untouched, so the line a
// pair is formed with is the last NON-SKIPPED line.  A line whose contig is in no
// avoid pair can never be skipped: chains of dependent decisions only run through
// consecutive "involved" lines and are resolved by walking back to the nearest
// line that is certainly kept and replaying forward (no limit on the length of the walk).
__device__ __forceinline__ bool avoid_has(const DevTable &t, int a, int b)
{
    if (a < 0 || b < 0 || a == b) return false;
    unsigned long long key = ((unsigned long long)(uint32_t)a << 32) | (uint32_t)b;
    int lo = 0, hi = t.navoid;
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (t.avoid[mid] < key) lo = mid + 1; else hi = mid;
    }
    return lo < t.navoid && t.avoid[lo] == key;
}

// copy the line that starts at abase coordinate st (length len) into the spare buffer and parse it
template <class Policy>
__device__ Rec parse_global_line(const ScanArgs &A, const Smem &S, long long st, int len)
{
    for (int i = 0; i < len; i++) S.s[SPARE_OFF + i] = A.abase[st + i];
    S.s[SPARE_OFF + len] = '\n';
    return parse_any<Policy>(A, S, SPARE_OFF, SPARE_OFF + len, false, false);
}

// Thread 0: was entry 0 (the line before the first owned line) skipped, and if so which
// earlier line is the effective "previous" of the first owned line?  (make_links.py:72-74: a skipped
// line does not become prev_line.)
//
// The answer depends on the contig of prev_line when the loop reaches entry 0.  It is known for certain
// behind a line whose contig is in no avoid pair (or the buffer's first line): round 2's first version
// walked back to the nearest such anchor however far that was -- a walk to the start of the file, for
// every tile, once every contig is in some avoid pair (a late iteration with a few dozen scaffolds).
// Now the walk goes back W lines (4, then x4 while the answer stays open) and replays forward with the
// SET of states prev_line can be in: before the window's first line it is unknown; behind it it is that
// line (kept) or one of the few contigs that line is avoid-paired with (skipped: an older line we have
// not seen); every further line either keeps a state (pair avoided) or replaces it by itself, so the set
// shrinks to one state within a few lines unless the text alternates between the two contigs of one
// avoid pair.  The answer is taken only when every state left gives the same one; a window that reaches
// an anchor or the start of the buffer is exact by construction, so the result never depends on W.
struct AvState {
    int id;                   // contig of prev_line
    long long st;             // abase coordinate of that line, -1: a line before the window (position unknown)
    int len;
};
constexpr int AV_STATES = 8;

template <class Policy>
__device__ void resolve_entry0(const ScanArgs &A, const Smem &S, long long sb)
{
    S.extinfo[0] = 0;
    S.extinfo[1] = -1;
    const int v0 = S.idv[0];
    const long long g0 = sb + (long long)S.lines[0];
    const long long mis = (long long)A.mis;
    if (!idv_inv(v0) || g0 <= mis) return;            // certainly kept (or the buffer's first line: always "previous")
    for (long long W = 4;; W *= 4) {
        // ---- back: W lines, or to an anchor ----
        long long cur = g0, first_st = g0;
        bool anchored = false;
        for (long long k = 0; k < W; k++) {
            long long e = cur - 1;                    // the '\n' that ends the previous line
            long long st = e;
            while (st > mis && A.abase[st - 1] != '\n' && e - st <= SALSA_MAX_LINE) st--;
            if (e - st > SALSA_MAX_LINE) {
                report(A.stats, A.base_offset + (uint64_t)(st - mis), SALSA_DATA_LINE_TOO_LONG);
                return;
            }
            first_st = st;
            Rec r = parse_global_line<Policy>(A, S, st, (int)(e - st));
            bool inv = r.id >= 0 && (A.t.meta[r.id].flags & CTG_INVOLVED);
            if (!inv || st <= mis) { anchored = true; break; }
            cur = st;
        }
        // ---- forward: the states prev_line can be in ----
        AvState set[AV_STATES];
        int ns = 0;
        bool open_ = false;                           // true: the window cannot settle the question
        long long pos = first_st;
        bool first = true;
        while (pos < g0 && !open_) {
            long long en = pos;
            while (en < g0 - 1 && A.abase[en] != '\n') en++;
            Rec r = parse_global_line<Policy>(A, S, pos, (int)(en - pos));     // (its length was checked on the way back)
            if (first) {
                first = false;
                set[ns++] = AvState{r.id, pos, (int)(en - pos)};             // the line itself: kept
                if (!anchored && r.id >= 0) {
                    // ... or skipped, behind an unseen line on a contig it is avoid-paired with
                    const unsigned long long k0 = (unsigned long long)(uint32_t)r.id << 32;
                    int lo = 0, hi = A.t.navoid;
                    while (lo < hi) {
                        int mid = (lo + hi) >> 1;
                        if (A.t.avoid[mid] < k0) lo = mid + 1; else hi = mid;
                    }
                    for (; lo < A.t.navoid && (A.t.avoid[lo] >> 32) == (unsigned long long)(uint32_t)r.id; lo++) {
                        if (ns == AV_STATES) { open_ = true; break; }
                        set[ns++] = AvState{(int)(uint32_t)A.t.avoid[lo], -1, 0};
                    }
                }
            } else {
                int m = 0;
                bool replaced = false;
                for (int q = 0; q < ns; q++) {
                    if (avoid_has(A.t, set[q].id, r.id)) set[m++] = set[q];   // pair avoided: prev_line stays
                    else replaced = true;
                }
                if (replaced) set[m++] = AvState{r.id, pos, (int)(en - pos)};
                ns = m;
            }
            pos = en + 1;
        }
        if (!open_) {
            // entry 0 against every state that is left
            int nskip = 0;
            for (int q = 0; q < ns; q++) nskip += avoid_has(A.t, set[q].id, idv_id(v0)) ? 1 : 0;
            if (nskip == 0) return;                   // kept, whatever came before
            if (nskip == ns && ns == 1 && set[0].st >= 0) {
                *S.ext = parse_global_line<Policy>(A, S, set[0].st, set[0].len);   // stays in the spare buffer for byte compares
                S.extinfo[0] = 1;
                S.extinfo[1] = set[0].id;
                return;
            }
        }
        // undecided: look further back (an anchored window is always decided: one state, position known)
    }
}

// Process the lines whose first byte lies in (lo, hi] (abase coordinates).
// sb: abase coordinate of smem byte 0.  WRITE_SP: also fill the special bitmap
// of the scanned chunks (the full-tile call; sub-range calls reuse it).
template <class Policy, bool WRITE_SP, bool WARP_AUTONOMOUS>
__device__ void process_range(const typename Policy::Args &PA, const ScanArgs &A, const Smem &S, long long sb, long long lo,
                              long long hi, long long l0, int sent, bool trunc, uint32_t &nlines_acc, uint32_t &npairs_acc)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint8_t *s = S.s;
    // ---- classify + newline scan: q in [lo, hi) -> line start q + 1 (valid iff q >= mis and q + 1 < dend)
    // 32 bytes (two 16-byte chunks) per lane and iteration: one ballot / rank / store per 32 bytes
    const int nchunks = (int)((hi - lo) >> 4);
    const int npair = (nchunks + 1) >> 1;
    const int ppw = (npair + NWARPS - 1) / NWARPS;
    int cnt = 0;                         // this warp's running count (uniform)
    const uint4 *s128 = reinterpret_cast<const uint4 *>(s);
    uint16_t *sp16 = reinterpret_cast<uint16_t *>(S.sp);
    // everything below in smem indices (int): a newline at index x counts iff vlo <= x < vhi
    const int c0 = (int)((lo - sb) >> 4);                    // first chunk of the range (even: ranges start on 32 bytes)
    const long long vlo_ll = (long long)A.mis - sb, vhi_ll = (long long)A.dend - 1 - sb;
    const int vlo = vlo_ll < 0 ? 0 : (int)vlo_ll;
    const int vhi = vhi_ll > (long long)SMEM_BYTES ? SMEM_BYTES : (int)vhi_ll;
    for (int it = 0; it < ppw; it += 32) {
        int pr = warp * ppw + it + lane;
        bool act = (it + lane) < ppw && pr < npair;
        int ci = c0 + 2 * pr;
        int x0 = ci << 4;                                    // smem index of the pair's byte 0
        uint32_t m32 = 0;
        if (act) {
            uint4 v0 = s128[ci];
            if (2 * pr + 1 < nchunks) {
                uint4 v1 = s128[ci + 1];
                m32 = nl16_of(v0) | (nl16_of(v1) << 16);
                if (WRITE_SP) S.sp[ci >> 1] = sp16_of(v0) | (sp16_of(v1) << 16);
            } else {                                         // odd tail: the data ends inside this pair
                m32 = nl16_of(v0);
                if (WRITE_SP) sp16[ci] = (uint16_t)sp16_of(v0);
            }
            // drop newlines outside the data or ending it (q + 1 < dend, q >= mis)
            if (x0 < vlo || x0 + 32 > vhi) {
                int a = vlo - x0, b = vhi - x0;
                a = a < 0 ? 0 : (a > 32 ? 32 : a);
                b = b < 0 ? 0 : (b > 32 ? 32 : b);
                uint32_t lowb = b >= 32 ? 0xFFFFFFFFu : ((1u << b) - 1u), lowa = a >= 32 ? 0xFFFFFFFFu : ((1u << a) - 1u);
                m32 &= b > a ? (lowb & ~lowa) : 0u;
            }
        }
        int sidx = x0 + 1;                     // smem index of the byte after the pair's byte 0
        unsigned has = __ballot_sync(0xffffffffu, m32 != 0u);
        unsigned multi = __ballot_sync(0xffffffffu, (m32 & (m32 - 1u)) != 0u);
        if (multi == 0u) {                     // at most one newline per 32 bytes: rank by ballot
            int pos = cnt + __popc(has & ((1u << lane) - 1u));
            if (m32 && pos < WCAP) S.seg[warp * WCAP + pos] = (uint16_t)(sidx + __ffs(m32) - 1);
            cnt += __popc(has);
        } else {
            int c = __popc(m32), inc = c;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                int o = __shfl_up_sync(0xffffffffu, inc, d);
                if (lane >= d) inc += o;
            }
            int pos = cnt + inc - c;
            int tot = __shfl_sync(0xffffffffu, inc, 31);
            while (m32) {
                int b = __ffs(m32) - 1;
                m32 &= m32 - 1;
                if (pos < WCAP) S.seg[warp * WCAP + pos] = (uint16_t)(sidx + b);
                pos++;
            }
            cnt += tot;
        }
    }
    if (lane == 0) {
        S.wcount[warp] = cnt;
        if (cnt > WCAP) atomicOr(S.flagp, 1);
    }
    __syncthreads();
    if (*S.flagp) return;                // only from the full-tile call; the caller falls back to subtiles
    if (WARP_AUTONOMOUS) {
        // Each warp parses and pairs the lines that START in its own 1/8 of the range, 32 per step.
        // No block barrier, no line-list compaction: the record of a step's last lane reaches the
        // next step by shuffle, and a warp's very first "previous" line is the last line start of
        // the nearest non-empty warp before it (or found by the backward search below).
        if (cnt == 0 && !Policy::LAZY) return;
        const uint16_t *my = S.seg + warp * WCAP;
        int prev_start = 0;
        if (cnt > 0) {
            int v = warp - 1;
            while (v >= 0 && S.wcount[v] == 0) v--;
            if (v >= 0) prev_start = S.seg[v * WCAP + S.wcount[v] - 1];
            else {
                long long lowpos = l0 > (long long)A.mis ? l0 : (long long)A.mis;
                int low = (int)(lowpos - sb);
                int first = my[0], found = -1;
                for (int base = first - 2; base >= low && found < 0; base -= 32) {
                    int idx = base - lane;
                    bool hit = idx >= low && s[idx] == '\n';
                    unsigned m = __ballot_sync(0xffffffffu, hit);
                    if (m) found = base - (__ffs(m) - 1);
                }
                if (found >= 0) prev_start = found + 1;
                else {
                    prev_start = low;                // first line of the buffer, or a line longer than the halo
                    if (lowpos != (long long)A.mis && lane == 0)
                        report(A.stats, A.base_offset + (uint64_t)(lowpos - A.mis), SALSA_DATA_LINE_TOO_LONG);
                }
            }
        }
        const int first_line_idx = (int)((long long)A.mis - sb);
        if (Policy::LAZY) {
            // Three phases.  Phase 1, every line: field layout only (line_shape) -- no name lookup, no digit
            // conversion, no read-name packing; a pair of plain lines is dropped here unless the policy wants it
            // (links: contig names differ and the read names may be equal).  What survives, and every pair with
            // a line that is not plain, is queued per warp.  Phase 2: whenever a warp has 32 waiting it runs the
            // ordinary path on them with all lanes busy (parse_any on both lines, Policy::pair).  Phase 3: what
            // is left in the eight queues at the end (< 32 each) is spread over the whole CTA, again with full
            // warps.  Results are identical by construction: phase 1 only drops pairs for which pair() neither
            // counts nor reports.
            uint8_t *q = S.queue + warp * QCAP;       // [0, 62) line indices, [64] leftover count, [66, 68) prev_start
            int qn = 0;
            const unsigned lt = (1u << lane) - 1u;
            auto full_pair = [&](const uint16_t *seg, int kk, int pst) {
                int ls = (int)seg[kk], ps = kk > 0 ? (int)seg[kk - 1] : pst;
                Rec rec = parse_any<Policy>(A, S, ls, sent, trunc, true);
                Rec pv = parse_any<Policy>(A, S, ps, sent, trunc, true);
                uint64_t off = A.base_offset + (uint64_t)((long long)ls + sb - (long long)A.mis);
                bool prev_is_first = sb <= 0 && ps == first_line_idx;
                if (Policy::pair(PA, A, s, pv, rec, off, prev_is_first)) npairs_acc++;
            };
            for (int k0 = -1; k0 + 1 < cnt; k0 += 31) {
                int k = k0 + lane;
                int lstart = 0;
                Shape h;
                h.w0 = h.w1 = h.w2 = h.w3 = 0; h.nl = NOT_PLAIN; h.rl = h.d0 = h.d1 = 0;
                if (k < cnt) {
                    lstart = k < 0 ? prev_start : (int)my[k];
                    h = line_shape(S, lstart, sent);
                }
                Shape ph;
                ph.w0 = __shfl_up_sync(0xffffffffu, h.w0, 1); ph.w1 = __shfl_up_sync(0xffffffffu, h.w1, 1);
                ph.w2 = __shfl_up_sync(0xffffffffu, h.w2, 1); ph.w3 = __shfl_up_sync(0xffffffffu, h.w3, 1);
                ph.nl = __shfl_up_sync(0xffffffffu, h.nl, 1); ph.rl = __shfl_up_sync(0xffffffffu, h.rl, 1);
                ph.d0 = __shfl_up_sync(0xffffffffu, h.d0, 1); ph.d1 = __shfl_up_sync(0xffffffffu, h.d1, 1);
                int pstart = __shfl_up_sync(0xffffffffu, lstart, 1);
                bool cand = false;
                if (lane >= 1 && k < cnt) {
                    nlines_acc++;
                    if (h.nl == NOT_PLAIN || ph.nl == NOT_PLAIN) cand = true;
                    else {
                        bool same = ((h.nl ^ ph.nl) | (h.w0 ^ ph.w0) | (h.w1 ^ ph.w1) | (h.w2 ^ ph.w2) | (h.w3 ^ ph.w3)) == 0u;
                        if (same && h.nl > 16u) same = bytes_equal_back(s, pstart + 16, lstart + 16, (int)h.nl - 16);
                        bool rmay = ((h.rl ^ ph.rl) | (h.d0 ^ ph.d0) | (h.d1 ^ ph.d1)) == 0u;
                        cand = Policy::lazy_candidate(same, rmay);
                    }
                }
                unsigned m = __ballot_sync(0xffffffffu, cand);
                if (cand) q[qn + __popc(m & lt)] = (uint8_t)k;
                qn += __popc(m);
                __syncwarp();
                if (qn >= 32) {
                    full_pair(my, (int)q[lane], prev_start);
                    int r = qn - 32;                         // < 31: move them to the front
                    uint8_t keep = lane < r ? q[32 + lane] : (uint8_t)0;
                    __syncwarp();
                    if (lane < r) q[lane] = keep;
                    qn = r;
                    __syncwarp();
                }
            }
            if (lane == 0) {
                q[64] = (uint8_t)qn;
                *reinterpret_cast<uint16_t *>(q + 66) = (uint16_t)prev_start;
            }
            __syncthreads();
            {
                // two lanes per pair (the odd one parses the previous line) so that twice as many warps share
                // the tail and the barrier above is held for half as long
                int total = 0;
#pragma unroll
                for (int w = 0; w < NWARPS; w++) total += S.queue[w * QCAP + 64];
                for (int base = 0; base < total; base += NTHREADS / 2) {
                    if (base + (warp << 4) >= total) break;              // warp-uniform: none of this warp's pairs exist
                    int t = base + (tid >> 1), w = 0, cw = S.queue[64];
                    const bool valid = t < total;
                    while (w < NWARPS - 1 && t >= cw) {
                        t -= cw;
                        w++;
                        cw = S.queue[w * QCAP + 64];
                    }
                    Rec rec;
                    rec.id = -1; rec.s_lo = rec.s_hi = rec.e_lo = rec.e_hi = 0; rec.name = rec.read = 0; rec.flags = 0;
#pragma unroll
                    for (int j = 0; j < RW; j++) rec.rw[j] = 0;
                    int ls = 0, ps = 0;
                    if (valid) {
                        const uint8_t *qw = S.queue + w * QCAP;
                        const uint16_t *seg = S.seg + w * WCAP;
                        int kk = qw[t];
                        ls = (int)seg[kk];
                        ps = kk > 0 ? (int)seg[kk - 1] : (int)*reinterpret_cast<const uint16_t *>(qw + 66);
                        rec = parse_any<Policy>(A, S, (lane & 1) ? ps : ls, sent, trunc, true);
                    }
                    Rec pv = rec_shfl_idx(rec, lane | 1);
                    if (valid && !(lane & 1)) {
                        uint64_t off = A.base_offset + (uint64_t)((long long)ls + sb - (long long)A.mis);
                        bool prev_is_first = sb <= 0 && ps == first_line_idx;
                        if (Policy::pair(PA, A, s, pv, rec, off, prev_is_first)) npairs_acc++;
                    }
                }
            }
            return;
        }
        if (cnt == 0) return;
        // steps overlap by one entry: lane 0 re-parses the previous step's last line and only plays
        // "previous" (1/32 more parsing, but no record has to live across steps)
        for (int k0 = -1; k0 + 1 < cnt; k0 += 31) {
            int k = k0 + lane;                       // entry -1 is the previous line, entries 0..cnt-1 are owned
            Rec rec;
            rec.id = -1; rec.s_lo = rec.s_hi = rec.e_lo = rec.e_hi = 0; rec.name = rec.read = 0; rec.flags = 0;
#pragma unroll
            for (int j = 0; j < RW; j++) rec.rw[j] = 0;
            int lstart = 0;
            if (k < cnt) {
                lstart = k < 0 ? prev_start : (int)my[k];
                rec = parse_any<Policy>(A, S, lstart, sent, trunc, true);
            }
            Rec pv = rec_shfl_up(rec);
            int pstart = __shfl_up_sync(0xffffffffu, lstart, 1);
            if (lane >= 1 && k < cnt) {
                nlines_acc++;
                uint64_t off = A.base_offset + (uint64_t)((long long)lstart + sb - (long long)A.mis);
                bool prev_is_first = sb <= 0 && pstart == first_line_idx;
                if (Policy::pair(PA, A, s, pv, rec, off, prev_is_first)) npairs_acc++;
            }
        }
        return;
    }
    int wb = 0, n = 0;
#pragma unroll
    for (int w = 0; w < NWARPS; w++) {
        int cw = S.wcount[w];
        if (w < warp) wb += cw;
        n += cw;
    }
    // lines[0] = the line before the first owned one; lines[1..n] = owned starts
    for (int i = lane; i < cnt; i += 32) S.lines[1 + wb + i] = S.seg[warp * WCAP + i];
    if (n > 0 && warp == 0) {
        // backward search for the start of the previous line: last '\n' before
        // lines[1] - 1, not below max(l0, mis)
        int first = 0;
        for (int w = 0; w < NWARPS; w++)
            if (S.wcount[w] > 0) { first = S.seg[w * WCAP]; break; }
        long long lowpos = l0 > (long long)A.mis ? l0 : (long long)A.mis;
        int low = (int)(lowpos - sb);
        int found = -1;
        for (int base = first - 2; base >= low && found < 0; base -= 32) {
            int idx = base - lane;
            bool hit = idx >= low && s[idx] == '\n';
            unsigned m = __ballot_sync(0xffffffffu, hit);
            if (m) found = base - (__ffs(m) - 1);
        }
        if (lane == 0) {
            if (found >= 0) S.lines[0] = (uint16_t)(found + 1);
            else if (lowpos == (long long)A.mis) S.lines[0] = (uint16_t)low;     // first line of the buffer
            else {
                S.lines[0] = (uint16_t)low;
                report(A.stats, A.base_offset + (uint64_t)(lowpos - A.mis), SALSA_DATA_LINE_TOO_LONG);
            }
        }
    }
    __syncthreads();
    if (n == 0) return;
    // ---- parse + pair ---------------------------------------------------
    const int K = n + 1;
    const int first_line_idx = (int)((long long)A.mis - sb);     // smem index of the buffer's first line (tile 0 only)
    bool av_fixed = false;       // avoid mode: the kept flags in S.idv are final (else: per-line replay below)
    if (Policy::AVOID) {
        // contig id (+ "in some avoid pair") of every line; thread 0 settles what happened before the tile
        for (int e = tid; e < K; e += NTHREADS) {
            Rec r = parse_any<Policy>(A, S, S.lines[e], sent, trunc, true);
            int v = r.id;
            if (v >= 0) v |= IDV_KEPT | ((A.t.meta[v].flags & CTG_INVOLVED) ? IDV_INVOLVED : 0);
            S.idv[e] = v;
        }
        __syncthreads();
        if (tid == 0) {
            resolve_entry0<Policy>(A, S, sb);
            if (S.extinfo[0] && S.idv[0] >= 0) S.idv[0] &= ~IDV_KEPT;
        }
        __syncthreads();
        // Which lines does the reference skip?  Line e is skipped iff its contig is avoid-paired with the contig of the
        // nearest line before it that was NOT skipped -- a left-to-right recurrence.  Its solution is the only fixed point
        // of "skipped(e) computed from the current guess of the lines before e", and iterating from "nothing is skipped"
        // gets every line right whose chain of skipped predecessors is shorter than the number of rounds: a handful in
        // any real BED.  Each round is O(1) per line, where replaying the chain from its start for every line
        // (round 1; kept below for texts that do not settle in 48 rounds) is quadratic in the length of a run of
        // avoid-listed lines -- 0.5 s per 10 GB once every scaffold is in some avoid pair.
        for (int round = 0; round < 48 && !av_fixed; round++) {
            int changed = 0;
            for (int e = 1 + tid; e < K; e += NTHREADS) {
                int v = S.idv[e];
                if (!idv_inv(v)) continue;                               // never skipped
                int j = e - 1;
                while (j > 0 && !idv_kept(S.idv[j])) j--;
                const int pid = (j == 0 && !idv_kept(S.idv[0])) ? S.extinfo[1] : idv_id(S.idv[j]);
                const bool kept = !avoid_has(A.t, pid, idv_id(v));
                changed |= (kept != ((v & IDV_KEPT) != 0));
                S.idv[e] = kept ? (v | IDV_NEXT) : (v & ~IDV_NEXT);      // (only this thread writes entry e; bit KEPT is untouched)
            }
            changed = __syncthreads_or(changed);
            for (int e = 1 + tid; e < K; e += NTHREADS) {
                int v = S.idv[e];
                if (idv_inv(v)) S.idv[e] = (v & IDV_NEXT) ? (v | IDV_KEPT) : (v & ~IDV_KEPT);
            }
            __syncthreads();
            av_fixed = changed == 0;
        }
    }
    for (int r0 = 0; r0 < K; r0 += NTHREADS) {
        int e = r0 + tid;
        Rec rec;
        rec.id = -1; rec.s_lo = rec.s_hi = rec.e_lo = rec.e_hi = 0; rec.name = rec.read = 0; rec.flags = 0;
#pragma unroll
        for (int j = 0; j < RW; j++) rec.rw[j] = 0;
        int lstart = 0;
        if (e < K) {
            lstart = S.lines[e];
            rec = parse_any<Policy>(A, S, lstart, sent, trunc, true);
        }
        if (lane == 31) S.carry[warp + 1] = rec;
        __syncthreads();
        Rec pv = rec_shfl_up(rec);
        int pstart = __shfl_up_sync(0xffffffffu, lstart, 1);
        if (lane == 0) { pv = S.carry[warp]; pstart = -2; }
        if (e >= 1 && e < K) {
            nlines_acc++;
            uint64_t off = A.base_offset + (uint64_t)((long long)lstart + sb - (long long)A.mis);
            bool skipped = false;
            if (Policy::AVOID) {
                // the line this one is paired with is the last non-skipped line (make_links.py:72-74)
                if (av_fixed) {
                    skipped = !idv_kept(S.idv[e]);
                    if (!skipped && !idv_kept(S.idv[e - 1])) {
                        int j = e - 1;
                        while (j > 0 && !idv_kept(S.idv[j])) j--;
                        if (j == 0 && !idv_kept(S.idv[0])) pv = *S.ext;
                        else pv = parse_any<Policy>(A, S, S.lines[j], sent, trunc, true);
                    }
                } else if (idv_inv(S.idv[e - 1])) {
                    int j = e - 1;
                    while (j > 0 && idv_inv(S.idv[j])) j--;
                    int prev = j, pid = idv_id(S.idv[j]);
                    if (j == 0 && S.extinfo[0]) { prev = -1; pid = S.extinfo[1]; }
                    for (int k = j + 1; k <= e; k++) {
                        int idk = idv_id(S.idv[k]);
                        bool sk = avoid_has(A.t, pid, idk);
                        if (k == e) { skipped = sk; break; }
                        if (!sk) { prev = k; pid = idk; }
                    }
                    if (!skipped && prev != e - 1) {
                        if (prev < 0) pv = *S.ext;
                        else pv = parse_any<Policy>(A, S, S.lines[prev], sent, trunc, true);
                    }
                }
            }
            // is the previous line the very first line of the caller's buffer?
            bool prev_is_first = sb <= 0 && ((e == 1) ? (S.lines[0] == first_line_idx) : (pstart == first_line_idx));
            if (!skipped && Policy::pair(PA, A, s, pv, rec, off, prev_is_first)) npairs_acc++;
        }
        __syncthreads();
        if (tid == NTHREADS - 1) S.carry[0] = rec;
    }
    __syncthreads();
}

constexpr size_t align16(size_t x) { return (x + 15) & ~(size_t)15; }
// Dynamic shared memory layout.  The plain kernels (warp-autonomous everywhere) need the staged
// bytes, the special bitmap, the per-warp line-start segments and the byte-mask table: 44.4 KB, so
// five CTAs fit one SM.  Avoid mode adds the spare line, the block-wide line list, the carry
// records and the per-line contig ids.
template <bool AV>
struct Layout {
    static constexpr size_t BYTES = AV ? (size_t)SMEM_BYTES : (size_t)SPARE_OFF;
    static constexpr size_t OFF_SP = align16(BYTES);
    static constexpr size_t OFF_SEG = OFF_SP + align16(SPWORDS * 4);
    static constexpr size_t OFF_MASK = OFF_SEG + align16(LCAP * 2);
    static constexpr size_t OFF_WCOUNT = OFF_MASK + align16(33 * RW * 4);
    static constexpr size_t OFF_QUEUE = OFF_WCOUNT + align16((NWARPS + 2) * sizeof(int));
    static constexpr size_t OFF_LINES = OFF_QUEUE + (AV ? 0 : NWARPS * QCAP);                   // avoid mode only from here
    static constexpr size_t OFF_CARRY = OFF_LINES + align16((LCAP + 8) * 2);
    static constexpr size_t OFF_IDV = OFF_CARRY + align16((NWARPS + 1) * sizeof(Rec));
    static constexpr size_t OFF_EXT = OFF_IDV + align16((LCAP + 8) * sizeof(int));
    static constexpr size_t OFF_EXTINFO = OFF_EXT + align16(sizeof(Rec));
    static constexpr size_t TOTAL = AV ? OFF_EXTINFO + 16 : OFF_LINES;
};
template <class Policy>
constexpr size_t dyn_smem() { return Layout<Policy::AVOID>::TOTAL; }

template <class Policy>
__global__ void __launch_bounds__(NTHREADS, Policy::AVOID ? 4 : SALSA_MINBLOCKS) bed_scan_kernel(typename Policy::Args PA, ScanArgs A)
{
    extern __shared__ __align__(128) uint8_t smem[];
    Smem S;
    typedef Layout<Policy::AVOID> LY;
    S.s = smem;
    S.sp = reinterpret_cast<uint32_t *>(smem + LY::OFF_SP);
    S.seg = reinterpret_cast<uint16_t *>(smem + LY::OFF_SEG);
    S.lowmask = reinterpret_cast<uint32_t (*)[RW]>(smem + LY::OFF_MASK);
    S.wcount = reinterpret_cast<int *>(smem + LY::OFF_WCOUNT);
    S.flagp = S.wcount + NWARPS;
    S.queue = smem + LY::OFF_QUEUE;
    S.lines =reinterpret_cast<uint16_t *>(smem + LY::OFF_LINES);      // the rest exists in avoid mode only
    S.carry = reinterpret_cast<Rec *>(smem + LY::OFF_CARRY);
    S.idv = reinterpret_cast<int *>(smem + LY::OFF_IDV);
    S.ext = reinterpret_cast<Rec *>(smem + LY::OFF_EXT);
    S.extinfo = reinterpret_cast<int *>(smem + LY::OFF_EXTINFO);
    uint8_t *s = S.s;
    __shared__ __align__(8) uint64_t bar;

    const int tid = threadIdx.x;
    const long long t0 = (long long)blockIdx.x * TILE, t1 = t0 + TILE;
    const long long sb = t0 - HALO;                            // abase coordinate of smem[0]
    const long long dend16 = ((long long)A.dend + 15) & ~15ll;
    const long long l0 = sb < 0 ? 0 : sb;
    const long long l1 = (t1 + HALO) < dend16 ? (t1 + HALO) : dend16;
    const uint32_t nb = (uint32_t)(l1 - l0);
    const int sent = (int)(l1 - sb);                           // smem index of the sentinel
    const bool trunc = l1 < (long long)A.dend;

    __shared__ uint32_t red[3];          // lines, counted pairs, warps done (epilogue)
    if (tid == 0) {
        *S.flagp = 0;
        red[0] = red[1] = red[2] = 0;
        uint32_t b = smem_u32(&bar);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0) {
        uint32_t b = smem_u32(&bar);
        uint32_t dst = smem_u32(s + (l0 - sb));
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(nb) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                     "l"(A.abase + l0), "r"(nb), "r"(b)
                     : "memory");
    }
    // while the copy is in flight: sentinel bytes, byte-mask table, bitmap guard words
    if (tid < 32) s[sent + tid] = '\n';
    for (int i = tid; i < 33 * RW; i += NTHREADS) {
        int e = i >> 3, j = i & 7;
        int nbv = e - 4 * j;
        S.lowmask[e][j] = nbv >= 4 ? 0xFFFFFFFFu : (nbv <= 0 ? 0u : ((1u << (8 * nbv)) - 1u));
    }
    // bitmap words at and beyond the end of the staged bytes read as "special" (the chunk
    // classification below overwrites the halves it owns)
    {
        int w0 = (sent >> 5) - 1 + tid;
        if (tid < 8 && w0 >= 0 && w0 < SPWORDS) S.sp[w0] = 0xFFFFFFFFu;
    }
    if (tid < 32) {                      // one warp waits for the bulk copy; the rest meet it at the barrier below
        uint32_t b = smem_u32(&bar);
        uint32_t done = 0;
        while (!done) {
            asm volatile(
                "{\n\t.reg .pred p;\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\t"
                "selp.u32 %0, 1, 0, p;\n\t}"
                : "=r"(done)
                : "r"(b)
                : "memory");
        }
    }
    // bytes outside the caller's data inside the first/last 16-byte granule: neutral '\n'
    if (l0 < (long long)A.mis) {
        if (tid < (int)A.mis) s[(int)(0 - sb) + tid] = '\n';
    }
    if ((long long)A.dend < l1) {
        int extra = (int)(l1 - (long long)A.dend);
        if (tid < extra) s[(int)((long long)A.dend - sb) + tid] = '\n';
    }
    __syncthreads();

    const long long hi = t1 < dend16 ? t1 : dend16;
    // special bitmap of the two halos (the owned range is classified inside process_range)
    {
        const uint4 *s128 = reinterpret_cast<const uint4 *>(s);
        uint16_t *sp16 = reinterpret_cast<uint16_t *>(S.sp);
        const int nback = (int)((t0 - l0) >> 4), nfwd = (int)((l1 - hi) >> 4);
        for (int i = tid; i < nback + nfwd; i += NTHREADS) {
            long long q0 = i < nback ? l0 + ((long long)i << 4) : hi + ((long long)(i - nback) << 4);
            int ci = (int)((q0 - sb) >> 4);
            sp16[ci] = (uint16_t)sp16_of(s128[ci]);
        }
    }
    uint32_t nlines = 0, npairs = 0;
    // avoid mode needs the block-wide line list (its walk-back crosses warp regions)
    process_range<Policy, true, !Policy::AVOID>(PA, A, S, sb, t0, hi, l0, sent, trunc, nlines, npairs);
    if (*S.flagp) {
        // > WCAP line starts inside one warp's 4 KB: redo the tile in 2 KB ranges,
        // which cannot overflow (a line start needs at least one byte)
        __syncthreads();
        nlines = npairs = 0;
        for (long long lo = t0; lo < hi; lo += SUBTILE) {
            __syncthreads();
            if (tid == 0) *S.flagp = 0;
            __syncthreads();
            long long h2 = lo + SUBTILE < hi ? lo + SUBTILE : hi;
            process_range<Policy, false, !Policy::AVOID>(PA, A, S, sb, lo, h2, l0, sent, trunc, nlines, npairs);
        }
    }
    // ---- statistics: one atomic pair per CTA -----------------------------
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        nlines += __shfl_xor_sync(0xffffffffu, nlines, d);
        npairs += __shfl_xor_sync(0xffffffffu, npairs, d);
    }
    // no barrier: every warp adds its sums in shared memory, the last one to do so owns the global update
    if ((tid & 31) == 0) {
        atomicAdd(&red[0], nlines);
        atomicAdd(&red[1], npairs);
        __threadfence_block();
        if (atomicAdd(&red[2], 1u) == NWARPS - 1) {
            uint32_t tl = atomicAdd(&red[0], 0u), tp = atomicAdd(&red[1], 0u);
            if (tl) atomicAdd((unsigned long long *)&A.stats[1], (unsigned long long)tl);
            if (tp) atomicAdd((unsigned long long *)&A.stats[2], (unsigned long long)tp);
        }
    }
}

template <class Policy>
int launch_scan(const typename Policy::Args &PA, const DevTable &t, const void *d_bed, size_t nbytes, uint64_t base_offset,
                uint64_t *stats, cudaStream_t st)
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
    uint64_t ntiles = (A.dend + TILE - 1) / TILE;
    if (ntiles > 0x7fffffffull) SALSA_FAIL(SALSA_ERR_ARG, "BED chunk too large (%zu bytes)", nbytes);
    // per device and cheap, so simply every launch
    CUDA_TRY(cudaFuncSetAttribute(bed_scan_kernel<Policy>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn_smem<Policy>()));
    bed_scan_kernel<Policy><<<(unsigned)ntiles, NTHREADS, dyn_smem<Policy>(), st>>>(PA, A);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}

}   // namespace bedscan
