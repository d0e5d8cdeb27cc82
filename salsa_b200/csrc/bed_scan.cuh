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
//   * warps find newlines with an exact SWAR zero-byte test, 16 bytes per lane,
//     and compact the line starts with a warp scan;
//   * one thread parses one line from shared memory (first four fields) and
//     hands its record to the next lane by warp shuffle -- the "previous line";
//   * Policy::pair() decides what an adjacent pair contributes.
#pragma once
#include "common.cuh"

namespace bedscan {

constexpr int TILE = 32768;
constexpr int HALO = SALSA_MAX_LINE;   // 1024
constexpr int NTHREADS = 256;
constexpr int NWARPS = NTHREADS / 32;
constexpr int LCAP = 2048;             // line starts per scan range
constexpr int WCAP = LCAP / NWARPS;    // per-warp segment
constexpr int SUBTILE = 2048;          // fallback range: <= 2048 line starts by construction
constexpr int SMEM_BYTES = HALO + TILE + HALO + 32;

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
    uint32_t read;        // roff | length up to the first '/' << 16
    uint32_t rfull;       // full length of the read token
    uint32_t flags;       // nf (bits 0-2) | numbad (bit 3) | toolong (bit 4)
};

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

// exact per-byte "== '\n'" mask: 0x80 in every byte of w that equals 0x0A
__device__ __forceinline__ uint32_t nl_mask(uint32_t w)
{
    uint32_t x = w ^ 0x0A0A0A0Au;
    uint32_t t = (x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu;
    return ~(t | x | 0x7F7F7F7Fu);
}

// 0x80-per-byte mask -> 4-bit mask (bit j = byte j)
__device__ __forceinline__ uint32_t bytes_to_nib(uint32_t z) { return (((z >> 7) * 0x00204081u) >> 21) & 15u; }

__device__ __forceinline__ void report(uint64_t *stats, uint64_t file_off, int code)
{
    atomicMin((unsigned long long *)&stats[0], (unsigned long long)((file_off << 8) | (uint64_t)code));
}

// ---- contig name -> id ------------------------------------------------
__device__ __forceinline__ int lookup_name(const DevTable &t, const uint8_t *s, int noff, int nlen)
{
    // pack the first 16 name bytes from aligned smem words
    const uint32_t *s32 = reinterpret_cast<const uint32_t *>(s);
    int a = noff >> 2;
    int sh = (noff & 3) * 8;
    uint32_t x0 = s32[a], x1 = s32[a + 1], x2 = s32[a + 2], x3 = s32[a + 3], x4 = s32[a + 4];
    uint32_t w0 = __funnelshift_r(x0, x1, sh), w1 = __funnelshift_r(x1, x2, sh);
    uint32_t w2 = __funnelshift_r(x2, x3, sh), w3 = __funnelshift_r(x3, x4, sh);
    if (nlen < 16) {
        int nb;
        nb = nlen;      w0 = nb >= 4 ? w0 : (nb <= 0 ? 0u : (w0 & ((1u << (8 * nb)) - 1u)));
        nb = nlen - 4;  w1 = nb >= 4 ? w1 : (nb <= 0 ? 0u : (w1 & ((1u << (8 * nb)) - 1u)));
        nb = nlen - 8;  w2 = nb >= 4 ? w2 : (nb <= 0 ? 0u : (w2 & ((1u << (8 * nb)) - 1u)));
        nb = nlen - 12; w3 = nb >= 4 ? w3 : (nb <= 0 ? 0u : (w3 & ((1u << (8 * nb)) - 1u)));
    }
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

// ---- one line -> record (first four whitespace-separated fields) -------
// s: smem bytes; p: smem index of the line's first byte; sent: smem index of
// the '\n' sentinel that ends the staged bytes; trunc: the sentinel stands for
// a halo cut, not for the end of the data.  Field rules are str.split()'s
// (make_links.py:62); numbers are strict [+-]?[0-9]{1,18}.
__device__ __forceinline__ Rec parse_line(const DevTable &t, const uint8_t *s, int p, int sent, bool trunc)
{
    Rec r;
    r.id = -1; r.s_lo = r.s_hi = r.e_lo = r.e_hi = 0; r.name = 0; r.read = 0; r.rfull = 0; r.flags = 0;
    uint32_t c = s[p];
    while (c != '\n' && is_ws(c)) c = s[++p];
    int nf = 0;
    uint32_t bad = 0;
    if (c != '\n') {
        int noff = p;
        do { c = s[++p]; } while (!is_ws(c));
        int nlen = p - noff;
        r.name = (uint32_t)noff | ((uint32_t)nlen << 16);
        r.id = lookup_name(t, s, noff, nlen);
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
                r.read = (uint32_t)roff | ((uint32_t)(p - roff) << 16);
                while (!is_ws(c)) c = s[++p];                  // rest of the token
                r.rfull = (uint32_t)(p - roff);
            }
        }
    }
    r.flags = (uint32_t)nf | bad | ((trunc && p >= sent) ? 16u : 0u);
    return r;
}

__device__ __forceinline__ bool bytes_equal_back(const uint8_t *s, int a, int b, int n)
{
    // compare from the end: consecutive read names differ in their last digits
    for (int i = n - 1; i >= 0; i--)
        if (s[a + i] != s[b + i]) return false;
    return true;
}

// contig equality of two adjacent records by NAME (ids when both are known)
__device__ __forceinline__ bool same_contig(const uint8_t *s, const Rec &pv, const Rec &cu)
{
    if (pv.id >= 0 && cu.id >= 0) return pv.id == cu.id;
    if ((pv.id ^ cu.id) < 0) return false;          // one known, one unknown
    int pl = pv.name >> 16, cl = cu.name >> 16;
    return pl == cl && bytes_equal_back(s, pv.name & 0xFFFF, cu.name & 0xFFFF, cl);
}

// Process the lines whose first byte lies in (lo, hi] (abase coordinates).
// sb: abase coordinate of smem byte 0.
template <class Policy>
__device__ void process_range(const typename Policy::Args &PA, const ScanArgs &A, uint8_t *s, uint16_t *seg, uint16_t *lines,
                              Rec *carry, int *wcount, int *flagp, long long sb, long long lo, long long hi, long long l0,
                              int sent, bool trunc, uint32_t &nlines_acc, uint32_t &npairs_acc)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // ---- newline scan: q in [lo, hi) -> line start q + 1 ---------------
    // valid iff q >= mis and q + 1 < dend
    const int nchunks = (int)((hi - lo) >> 4);
    const int cpw = (nchunks + NWARPS - 1) / NWARPS;
    int cnt = 0;                         // this warp's running count (uniform)
    const uint4 *s128 = reinterpret_cast<const uint4 *>(s);
    for (int it = 0; it < cpw; it += 32) {
        int ch = warp * cpw + it + lane;
        bool act = (it + lane) < cpw && ch < nchunks;
        long long q0 = lo + ((long long)ch << 4);
        uint32_t m16 = 0;
        if (act) {
            uint4 v = s128[(q0 - sb) >> 4];
            m16 = bytes_to_nib(nl_mask(v.x)) | (bytes_to_nib(nl_mask(v.y)) << 4) | (bytes_to_nib(nl_mask(v.z)) << 8) |
                  (bytes_to_nib(nl_mask(v.w)) << 12);
            // drop newlines outside the data or ending it (q + 1 < dend, q >= mis)
            if (q0 + 16 >= (long long)A.dend || q0 < (long long)A.mis) {
                long long klo = (long long)A.mis - q0, khi = (long long)A.dend - 1 - q0;
                int a = klo < 0 ? 0 : (klo > 16 ? 16 : (int)klo);
                int b = khi < 0 ? 0 : (khi > 16 ? 16 : (int)khi);
                m16 &= b > a ? (((1u << b) - 1u) & ~((1u << a) - 1u)) : 0u;
            }
        }
        int c = __popc(m16);
        int inc = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += o;
        }
        int pos = cnt + inc - c;
        int tot = __shfl_sync(0xffffffffu, inc, 31);
        int sidx = (int)(q0 - sb) + 1;         // smem index of the byte after chunk byte 0
        while (m16) {
            int b = __ffs(m16) - 1;
            m16 &= m16 - 1;
            if (pos < WCAP) seg[warp * WCAP + pos] = (uint16_t)(sidx + b);
            pos++;
        }
        cnt += tot;
    }
    if (lane == 0) {
        wcount[warp] = cnt;
        if (cnt > WCAP) atomicOr(flagp, 1);
    }
    __syncthreads();
    if (*flagp) return;                  // only from the full-tile call; the caller falls back to subtiles
    int wb = 0, n = 0;
#pragma unroll
    for (int w = 0; w < NWARPS; w++) {
        int cw = wcount[w];
        if (w < warp) wb += cw;
        n += cw;
    }
    // lines[0] = the line before the first owned one; lines[1..n] = owned starts
    for (int i = lane; i < cnt; i += 32) lines[1 + wb + i] = seg[warp * WCAP + i];
    if (n > 0 && warp == 0) {
        // backward search for the start of the previous line: last '\n' before
        // lines[1] - 1, not below max(l0, mis)
        int first = 0;
        for (int w = 0; w < NWARPS; w++)
            if (wcount[w] > 0) { first = seg[w * WCAP]; break; }
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
            if (found >= 0) lines[0] = (uint16_t)(found + 1);
            else if (lowpos == (long long)A.mis) lines[0] = (uint16_t)low;     // first line of the buffer
            else {
                lines[0] = (uint16_t)low;
                report(A.stats, A.base_offset + (uint64_t)(lowpos - A.mis), SALSA_DATA_LINE_TOO_LONG);
            }
        }
    }
    __syncthreads();
    if (n == 0) return;
    // ---- parse + pair ---------------------------------------------------
    const int K = n + 1;
    const int first_line_idx = (int)((long long)A.mis - sb);     // smem index of the buffer's first line (tile 0 only)
    for (int r0 = 0; r0 < K; r0 += NTHREADS) {
        int e = r0 + tid;
        Rec rec;
        rec.id = -1; rec.s_lo = rec.s_hi = rec.e_lo = rec.e_hi = 0; rec.name = rec.read = rec.rfull = 0; rec.flags = 0;
        int lstart = 0;
        if (e < K) {
            lstart = lines[e];
            rec = parse_line(A.t, s, lstart, sent, trunc);
        }
        if (lane == 31) carry[warp + 1] = rec;
        __syncthreads();
        Rec pv;
        pv.id = __shfl_up_sync(0xffffffffu, rec.id, 1);
        pv.s_lo = __shfl_up_sync(0xffffffffu, rec.s_lo, 1);
        pv.s_hi = __shfl_up_sync(0xffffffffu, rec.s_hi, 1);
        pv.e_lo = __shfl_up_sync(0xffffffffu, rec.e_lo, 1);
        pv.e_hi = __shfl_up_sync(0xffffffffu, rec.e_hi, 1);
        pv.name = __shfl_up_sync(0xffffffffu, rec.name, 1);
        pv.read = __shfl_up_sync(0xffffffffu, rec.read, 1);
        pv.rfull = __shfl_up_sync(0xffffffffu, rec.rfull, 1);
        pv.flags = __shfl_up_sync(0xffffffffu, rec.flags, 1);
        int pstart = __shfl_up_sync(0xffffffffu, lstart, 1);
        if (lane == 0) { pv = carry[warp]; pstart = -2; }
        if (e >= 1 && e < K) {
            nlines_acc++;
            uint64_t off = A.base_offset + (uint64_t)((long long)lstart + sb - (long long)A.mis);
            // is the previous line the very first line of the caller's buffer?
            bool prev_is_first = (e == 1) ? (lines[0] == first_line_idx && sb <= 0) : (pstart == first_line_idx && sb <= 0);
            if (Policy::pair(PA, A, s, pv, rec, off, prev_is_first)) npairs_acc++;
        }
        __syncthreads();
        if (tid == NTHREADS - 1) carry[0] = rec;
    }
    __syncthreads();
}

template <class Policy>
__global__ void __launch_bounds__(NTHREADS, 4) bed_scan_kernel(typename Policy::Args PA, ScanArgs A)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *s = smem;                                         // SMEM_BYTES
    uint16_t *seg = reinterpret_cast<uint16_t *>(smem + SMEM_BYTES);        // LCAP
    uint16_t *lines = seg + LCAP;                              // LCAP + 8
    Rec *carry = reinterpret_cast<Rec *>(lines + LCAP + 8);    // NWARPS + 1
    int *wcount = reinterpret_cast<int *>(carry + NWARPS + 1); // NWARPS
    int *flagp = wcount + NWARPS;
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

    if (tid == 0) {
        *flagp = 0;
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
    // sentinel + pad (generic-proxy writes to bytes the bulk copy does not touch)
    if (tid < 32) s[sent + tid] = '\n';
    {
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

    uint32_t nlines = 0, npairs = 0;
    long long hi = t1 < dend16 ? t1 : dend16;
    process_range<Policy>(PA, A, s, seg, lines, carry, wcount, flagp, sb, t0, hi, l0, sent, trunc, nlines, npairs);
    if (*flagp) {
        // > WCAP line starts inside one warp's 4 KB: redo the tile in 2 KB ranges,
        // which cannot overflow (a line start needs at least one byte)
        __syncthreads();
        nlines = npairs = 0;
        for (long long lo = t0; lo < hi; lo += SUBTILE) {
            __syncthreads();
            if (tid == 0) *flagp = 0;
            __syncthreads();
            long long h2 = lo + SUBTILE < hi ? lo + SUBTILE : hi;
            process_range<Policy>(PA, A, s, seg, lines, carry, wcount, flagp, sb, lo, h2, l0, sent, trunc, nlines, npairs);
        }
    }
    // ---- statistics: one atomic pair per CTA -----------------------------
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        nlines += __shfl_xor_sync(0xffffffffu, nlines, d);
        npairs += __shfl_xor_sync(0xffffffffu, npairs, d);
    }
    __shared__ uint32_t red[2];
    if (tid == 0) red[0] = red[1] = 0;
    __syncthreads();
    if ((tid & 31) == 0) {
        atomicAdd(&red[0], nlines);
        atomicAdd(&red[1], npairs);
    }
    __syncthreads();
    if (tid == 0) {
        if (red[0]) atomicAdd((unsigned long long *)&A.stats[1], (unsigned long long)red[0]);
        if (red[1]) atomicAdd((unsigned long long *)&A.stats[2], (unsigned long long)red[1]);
    }
}

constexpr size_t DYN_SMEM = SMEM_BYTES + LCAP * 2 + (LCAP + 8) * 2 + (NWARPS + 1) * sizeof(Rec) + (NWARPS + 1) * sizeof(int);

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
    CUDA_TRY(cudaFuncSetAttribute(bed_scan_kernel<Policy>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DYN_SMEM));
    bed_scan_kernel<Policy><<<(unsigned)ntiles, NTHREADS, DYN_SMEM, st>>>(PA, A);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}

}   // namespace bedscan
