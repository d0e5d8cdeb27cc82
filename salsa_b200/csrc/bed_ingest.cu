// bed_ingest.cu -- BED text -> contig-end link histogram, one pass over HBM.
//
// Replaces the per-line loop of the reference's make_links.py:60-146
// (str.split, adjacent-line pairing, read.split('/')[0] test, B/E end
// classification, contig_links[key] += 1).
//
// Shape of the kernel (DESIGN.md has the long version):
//   * the BED is cut into TILE-byte tiles; one CTA stages its tile plus a
//     SALSA_MAX_LINE halo on either side in shared memory with a single 1-D
//     TMA bulk copy (cp.async.bulk ... mbarrier::complete_tx);
//   * a tile OWNS the lines whose first byte lies in (t0, t1]; it evaluates the
//     adjacent-line pairs whose second line it owns, so every pair is
//     evaluated exactly once and the line at offset 0 only ever plays
//     "previous" (make_links.py:66-68);
//   * warps find newlines with an exact SWAR zero-byte test, 16 bytes per lane,
//     and compact the line starts with a warp scan;
//   * one thread parses one line from shared memory (first four fields only),
//     looks the contig name up in the device name table (exact match, names
//     <= 16 bytes from the slot itself) and hands its record to the next lane
//     by warp shuffle -- the "previous line" of the reference;
//   * a counted pair is one 64-bit RED into the dense table (or one insert
//     into the sparse hash table).
// HBM traffic = the BED bytes once; everything else lives in smem / L2.
#include "common.cuh"

namespace {

constexpr int TILE = 32768;
constexpr int HALO = SALSA_MAX_LINE;   // 1024
constexpr int NTHREADS = 256;
constexpr int NWARPS = NTHREADS / 32;
constexpr int LCAP = 2048;             // line starts per scan range
constexpr int WCAP = LCAP / NWARPS;    // per-warp segment
constexpr int SUBTILE = 2048;          // fallback range: <= 2048 lines by construction
constexpr int SMEM_BYTES = HALO + TILE + HALO + 32;

enum { MODE_DENSE = 0, MODE_SPARSE = 1 };

struct IngestArgs {
    const uint8_t *abase;     // 16-byte aligned base (<= caller pointer)
    uint32_t mis;             // caller pointer - abase  (0..15)
    uint64_t dend;            // mis + nbytes : one past the last valid byte (abase coordinates)
    uint64_t base_offset;     // file offset of the caller's byte 0 (error reports)
    DevTable t;
    uint64_t *table;          // dense: counts; sparse: counts
    uint64_t *keys;           // sparse only
    uint64_t cap_mask;        // sparse only
    uint64_t *stats;
};

struct Rec {
    int id;            // contig id, -1 unknown
    int sum_lo, sum_hi;   // start + end (int64)
    uint32_t name;     // noff | nlen << 16   (smem offsets)
    uint32_t read;     // roff | rlen << 16
    uint32_t flags;    // nf (bits 0-2) | numbad (bit 3) | toolong (bit 4)
};

__device__ __forceinline__ bool is_ws(uint32_t c) { return c == 32u || (c - 9u) <= 4u; }

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// exact per-byte "== '\n'" mask: 0x80 in every byte of w that equals 0x0A
__device__ __forceinline__ uint32_t nl_mask(uint32_t w)
{
    uint32_t x = w ^ 0x0A0A0A0Au;
    uint32_t t = (x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu;
    return ~(t | x | 0x7F7F7F7Fu);
}

// 4-bit mask -> 0x80 in the matching bytes
__device__ __forceinline__ uint32_t nib_to_bytes(uint32_t nib)
{
    return ((nib & 1u) << 7) | ((nib & 2u) << 14) | ((nib & 4u) << 21) | ((nib & 8u) << 28);
}

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
// a halo cut, not for the end of the data.
__device__ __forceinline__ Rec parse_line(const DevTable &t, const uint8_t *s, int p, int sent, bool trunc)
{
    Rec r;
    r.id = -1; r.sum_lo = 0; r.sum_hi = 0; r.name = 0; r.read = 0; r.flags = 0;
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
        long long sum = 0;
#pragma unroll 1
        for (int f = 0; f < 2; f++) {
            while (c != '\n' && is_ws(c)) c = s[++p];
            if (c == '\n') break;
            nf++;
            bool neg = (c == '-');
            if (c == '-' || c == '+') c = s[++p];
            unsigned long long v = 0;
            int nd = 0;
            uint32_t d = c - '0';
            while (d <= 9u) { v = v * 10ull + d; nd++; c = s[++p]; d = c - '0'; }
            if (nd == 0 || nd > 18 || !is_ws(c)) bad = 8u;
            while (!is_ws(c)) c = s[++p];
            sum += neg ? -(long long)v : (long long)v;
        }
        r.sum_lo = (int)(uint32_t)(unsigned long long)sum;
        r.sum_hi = (int)(uint32_t)((unsigned long long)sum >> 32);
        if (nf == 3) {
            while (c != '\n' && is_ws(c)) c = s[++p];
            if (c != '\n') {
                nf = 4;
                int roff = p;
                while (c != '/' && !is_ws(c)) c = s[++p];      // read.split('/')[0]
                r.read = (uint32_t)roff | ((uint32_t)(p - roff) << 16);
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

template <int MODE>
__device__ __forceinline__ void emit(const IngestArgs &A, uint64_t key, uint64_t line_off)
{
    if (MODE == MODE_DENSE) {
        atomicAdd((unsigned long long *)&A.table[key], 1ull);
    } else {
        uint64_t slot = salsa_mix64(key) & A.cap_mask;
        for (uint64_t probe = 0; probe <= A.cap_mask; probe++) {
            unsigned long long k = ((volatile unsigned long long *)A.keys)[slot];
            if (k == ~0ull) {
                unsigned long long old = atomicCAS((unsigned long long *)&A.keys[slot], ~0ull, (unsigned long long)key);
                if (old == ~0ull || old == key) k = key;
                else k = old;
            }
            if (k == key) {
                atomicAdd((unsigned long long *)&A.table[slot], 1ull);
                return;
            }
            slot = (slot + 1) & A.cap_mask;
        }
        report(A.stats, line_off, SALSA_DATA_TABLE_FULL);
    }
}

// the reference's body for one adjacent-line pair: make_links.py:70-144
template <int MODE>
__device__ __forceinline__ bool eval_pair(const IngestArgs &A, const uint8_t *s, const Rec &pv, const Rec &cu,
                                          uint64_t line_off)
{
    uint32_t pnf = pv.flags & 7u, cnf = cu.flags & 7u;
    if ((pv.flags | cu.flags) & 16u) { report(A.stats, line_off, SALSA_DATA_LINE_TOO_LONG); return false; }
    if (pnf < 1 || cnf < 1) { report(A.stats, line_off, SALSA_DATA_INDEX_ERROR); return false; }   // :71
    bool same;
    if (pv.id >= 0 && cu.id >= 0) same = pv.id == cu.id;
    else if ((pv.id ^ cu.id) < 0) same = false;          // one known, one unknown
    else {
        int pl = pv.name >> 16, cl = cu.name >> 16;
        same = pl == cl && bytes_equal_back(s, pv.name & 0xFFFF, cu.name & 0xFFFF, cl);
    }
    if (same) return false;
    if (pnf < 4 || cnf < 4) { report(A.stats, line_off, SALSA_DATA_INDEX_ERROR); return false; }   // :76-77
    int pl = pv.read >> 16, cl = cu.read >> 16;
    if (pl != cl || !bytes_equal_back(s, pv.read & 0xFFFF, cu.read & 0xFFFF, cl)) return false;    // :78
    if ((pv.flags | cu.flags) & 8u) { report(A.stats, line_off, SALSA_DATA_VALUE_ERROR); return false; }  // :81-82
    if (pv.id < 0 || cu.id < 0) { report(A.stats, line_off, SALSA_DATA_KEY_ERROR); return false; }  // :101-102
    ContigMeta mp = A.t.meta[pv.id], mc = A.t.meta[cu.id];
    if (!(mp.flags & mc.flags & SALSA_CTG_HAS_LEN)) { report(A.stats, line_off, SALSA_DATA_KEY_ERROR); return false; }
    if (mp.len_floor == 0 || mc.len_floor == 0) {
        // float(len) == 0 raises ZeroDivisionError (:106-107); 0 < len < 1 floors to 0
        // as well but such lengths are rejected on the host.
        report(A.stats, line_off, SALSA_DATA_ZERO_DIV);
        return false;
    }
    if (!(mp.flags & mc.flags & SALSA_CTG_HAS_RE)) { report(A.stats, line_off, SALSA_DATA_KEY_ERROR); return false; }  // :113
    long long sp = (long long)(((unsigned long long)(uint32_t)pv.sum_hi << 32) | (uint32_t)pv.sum_lo);
    long long sc = (long long)(((unsigned long long)(uint32_t)cu.sum_hi << 32) | (uint32_t)cu.sum_lo);
    // B <=> pos <= len/2 <=> start+end <= floor(len); E otherwise (:108-139)
    uint32_t ep = sp > mp.len_floor, ec = sc > mc.len_floor;
    uint32_t first, second, ends;
    if (pv.id < cu.id) { first = pv.id; second = cu.id; ends = ep * 2 + ec; }     // :90-96
    else { first = cu.id; second = pv.id; ends = ec * 2 + ep; }
    uint64_t key = (((uint64_t)first * (uint64_t)A.t.n + second) << 2) | ends;
    emit<MODE>(A, key, line_off);
    return true;
}

// Process the lines whose first byte lies in (lo, hi] (abase coordinates).
// sb: abase coordinate of smem byte 0.
template <int MODE>
__device__ void process_range(const IngestArgs &A, uint8_t *s, uint16_t *seg, uint16_t *lines, Rec *carry,
                              int *wcount, int *flagp, long long sb, long long lo, long long hi, long long l0,
                              int sent, bool trunc, bool allow_overflow, uint32_t &nlines_acc, uint32_t &npairs_acc)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // ---- newline scan: q in [lo, hi) -> line start q + 1 ---------------
    // valid iff q >= mis (true for all q >= lo >= 0 except in tile 0) and q + 1 < dend
    const int nchunks = (int)((hi - lo) >> 4);
    const int cpw = (nchunks + NWARPS - 1) / NWARPS;
    int cnt = 0;                         // this warp's running count (uniform)
    const uint4 *s128 = reinterpret_cast<const uint4 *>(s);
    for (int it = 0; it < cpw; it += 32) {
        int ch = warp * cpw + it + lane;
        bool act = (it + lane) < cpw && ch < nchunks;
        long long q0 = lo + ((long long)ch << 4);
        uint32_t z0 = 0, z1 = 0, z2 = 0, z3 = 0;
        if (act) {
            uint4 v = s128[(q0 - sb) >> 4];
            z0 = nl_mask(v.x); z1 = nl_mask(v.y); z2 = nl_mask(v.z); z3 = nl_mask(v.w);
            // drop newlines outside the data or ending it (q + 1 < dend, q >= mis)
            if (q0 + 16 >= (long long)A.dend || q0 < (long long)A.mis) {
                // valid byte indices k: mis <= q0 + k <= dend - 2
                long long klo = (long long)A.mis - q0, khi = (long long)A.dend - 1 - q0;
                int a = klo < 0 ? 0 : (klo > 16 ? 16 : (int)klo);
                int b = khi < 0 ? 0 : (khi > 16 ? 16 : (int)khi);
                uint32_t vm = b > a ? (((1u << b) - 1u) & ~((1u << a) - 1u)) : 0u;
                z0 &= nib_to_bytes(vm & 15u);
                z1 &= nib_to_bytes((vm >> 4) & 15u);
                z2 &= nib_to_bytes((vm >> 8) & 15u);
                z3 &= nib_to_bytes((vm >> 12) & 15u);
            }
        }
        int c = __popc(z0) + __popc(z1) + __popc(z2) + __popc(z3);
        int inc = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += o;
        }
        int pos = cnt + inc - c;
        int tot = __shfl_sync(0xffffffffu, inc, 31);
        if (c) {
            int sidx = (int)(q0 - sb) + 1;     // smem index of the byte after chunk byte 0
            uint32_t zs[4] = {z0, z1, z2, z3};
#pragma unroll
            for (int k = 0; k < 4; k++) {
                uint32_t z = zs[k];
                while (z) {
                    int b = (__ffs(z) - 1) >> 3;
                    z &= z - 1;
                    if (pos < WCAP) seg[warp * WCAP + pos] = (uint16_t)(sidx + k * 4 + b);
                    pos++;
                }
            }
        }
        cnt += tot;
    }
    if (lane == 0) {
        wcount[warp] = cnt;
        if (cnt > WCAP) atomicOr(flagp, 1);
    }
    __syncthreads();
    int overflow = *flagp;
    if (overflow) {
        // only reachable from the full-tile call; the caller falls back to subtiles
        return;
    }
    (void)allow_overflow;
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
        int first;
        if (wcount[0] > 0) first = seg[0];
        else {
            first = 0;
            for (int w = 1; w < NWARPS; w++)
                if (wcount[w] > 0) { first = seg[w * WCAP]; break; }
        }
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
    for (int r0 = 0; r0 < K; r0 += NTHREADS) {
        int e = r0 + tid;
        Rec rec;
        rec.id = -1; rec.sum_lo = rec.sum_hi = 0; rec.name = rec.read = 0; rec.flags = 0;
        int lstart = 0;
        if (e < K) {
            lstart = lines[e];
            rec = parse_line(A.t, s, lstart, sent, trunc);
        }
        if (lane == 31) carry[warp + 1] = rec;
        __syncthreads();
        Rec pv;
        pv.id = __shfl_up_sync(0xffffffffu, rec.id, 1);
        pv.sum_lo = __shfl_up_sync(0xffffffffu, rec.sum_lo, 1);
        pv.sum_hi = __shfl_up_sync(0xffffffffu, rec.sum_hi, 1);
        pv.name = __shfl_up_sync(0xffffffffu, rec.name, 1);
        pv.read = __shfl_up_sync(0xffffffffu, rec.read, 1);
        pv.flags = __shfl_up_sync(0xffffffffu, rec.flags, 1);
        if (lane == 0) pv = carry[warp];
        if (e >= 1 && e < K) {
            nlines_acc++;
            uint64_t off = A.base_offset + (uint64_t)((long long)lstart + sb - (long long)A.mis);
            if (eval_pair<MODE>(A, s, pv, rec, off)) npairs_acc++;
        }
        __syncthreads();
        if (tid == NTHREADS - 1) carry[0] = rec;
    }
    __syncthreads();
}

template <int MODE>
__global__ void __launch_bounds__(NTHREADS, 4) bed_ingest_kernel(IngestArgs A)
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
    // sentinel + pad (written by generic proxy to bytes the bulk copy does not touch)
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
    long long hi = t1 < (long long)A.dend ? t1 : (((long long)A.dend + 15) & ~15ll);
    if (hi > t1) hi = t1;
    process_range<MODE>(A, s, seg, lines, carry, wcount, flagp, sb, t0, hi, l0, sent, trunc, true, nlines, npairs);
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
            process_range<MODE>(A, s, seg, lines, carry, wcount, flagp, sb, lo, h2, l0, sent, trunc, false, nlines, npairs);
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

template <int MODE>
int launch_ingest(salsa_ctx *ctx, const void *d_bed, size_t nbytes, uint64_t base_offset, uint64_t *table, uint64_t *keys,
                  size_t capacity, uint64_t *stats, cudaStream_t st)
{
    if (nbytes == 0) return SALSA_OK;
    IngestArgs A;
    uintptr_t p = (uintptr_t)d_bed;
    A.mis = (uint32_t)(p & 15u);
    A.abase = (const uint8_t *)(p - A.mis);
    A.dend = (uint64_t)A.mis + nbytes;
    A.base_offset = base_offset;
    A.t = ctx->t;
    A.table = table;
    A.keys = keys;
    A.cap_mask = capacity ? capacity - 1 : 0;
    A.stats = stats;
    uint64_t ntiles = (A.dend + TILE - 1) / TILE;
    if (ntiles > 0x7fffffffull) SALSA_FAIL(SALSA_ERR_ARG, "BED chunk too large (%zu bytes)", nbytes);
    static bool attr_set[2] = {false, false};
    if (!attr_set[MODE]) {
        CUDA_TRY(cudaFuncSetAttribute(bed_ingest_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DYN_SMEM));
        attr_set[MODE] = true;
    }
    bed_ingest_kernel<MODE><<<(unsigned)ntiles, NTHREADS, DYN_SMEM, st>>>(A);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}

__global__ void stats_reset_kernel(uint64_t *s)
{
    s[0] = ~0ull;
    s[1] = s[2] = s[3] = 0;
}

__global__ void sparse_reset_kernel(uint64_t *keys, uint64_t *counts, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) { keys[i] = ~0ull; counts[i] = 0; }
}

}   // namespace

extern "C" int salsa_ingest_stats_reset(uint64_t *d_stats, void *stream)
{
    if (!d_stats) SALSA_FAIL(SALSA_ERR_ARG, "d_stats is NULL");
    stats_reset_kernel<<<1, 1, 0, as_stream(stream)>>>(d_stats);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}

extern "C" int salsa_ingest_stats_read(const uint64_t *d_stats, salsa_ingest_stats *out, void *stream)
{
    if (!d_stats || !out) SALSA_FAIL(SALSA_ERR_ARG, "bad argument");
    uint64_t h[4];
    CUDA_TRY(cudaMemcpyAsync(h, d_stats, sizeof h, cudaMemcpyDeviceToHost, as_stream(stream)));
    CUDA_TRY(cudaStreamSynchronize(as_stream(stream)));
    if (h[0] == ~0ull) { out->err_code = 0; out->err_offset = 0; }
    else { out->err_code = (int32_t)(h[0] & 0xff); out->err_offset = h[0] >> 8; }
    out->lines = h[1];
    out->pairs = h[2];
    return SALSA_OK;
}

static int check_ingest_args(salsa_ctx *ctx, const void *d_bed, size_t nbytes, const void *table, const void *stats)
{
    if (!ctx || (!d_bed && nbytes) || !table || !stats) SALSA_FAIL(SALSA_ERR_ARG, "salsa_bed_ingest: NULL argument");
    if (ctx->n <= 0) SALSA_FAIL(SALSA_ERR_ARG, "salsa_bed_ingest: call salsa_names_build first");
    if (ctx->t.navoid > 0)
        SALSA_FAIL(SALSA_ERR_UNSUPPORTED, "salsa_bed_ingest: avoid set present; use salsa_bed_ingest_avoid");
    CUDA_TRY(cudaSetDevice(ctx->device));
    return SALSA_OK;
}

extern "C" int salsa_bed_ingest_dense(salsa_ctx *ctx, const void *d_bed, size_t nbytes, uint64_t base_offset,
                                      uint64_t *d_table, uint64_t *d_stats, void *stream)
{
    int rc = check_ingest_args(ctx, d_bed, nbytes, d_table, d_stats);
    if (rc) return rc;
    return launch_ingest<MODE_DENSE>(ctx, d_bed, nbytes, base_offset, d_table, NULL, 0, d_stats, as_stream(stream));
}

extern "C" int salsa_sparse_reset(uint64_t *d_keys, uint64_t *d_counts, size_t capacity, void *stream)
{
    if (!d_keys || !d_counts || capacity == 0 || (capacity & (capacity - 1)))
        SALSA_FAIL(SALSA_ERR_ARG, "salsa_sparse_reset: capacity must be a power of two");
    sparse_reset_kernel<<<148 * 8, 256, 0, as_stream(stream)>>>(d_keys, d_counts, capacity);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}

extern "C" int salsa_bed_ingest_sparse(salsa_ctx *ctx, const void *d_bed, size_t nbytes, uint64_t base_offset,
                                       uint64_t *d_keys, uint64_t *d_counts, size_t capacity, uint64_t *d_stats,
                                       void *stream)
{
    int rc = check_ingest_args(ctx, d_bed, nbytes, d_counts, d_stats);
    if (rc) return rc;
    if (!d_keys || capacity == 0 || (capacity & (capacity - 1)))
        SALSA_FAIL(SALSA_ERR_ARG, "salsa_bed_ingest_sparse: capacity must be a power of two");
    return launch_ingest<MODE_SPARSE>(ctx, d_bed, nbytes, base_offset, d_counts, d_keys, capacity, d_stats, as_stream(stream));
}

// Host-resident BED: chunked H2D copies overlapped with the kernel.  Chunks
// start on line starts and overlap by exactly one line (see salsa_b200.h).
extern "C" int salsa_bed_ingest_dense_host(salsa_ctx *ctx, const void *h_bed, size_t nbytes, void *d_stage,
                                           size_t chunk_bytes, uint64_t *d_table, uint64_t *d_stats)
{
    int rc = check_ingest_args(ctx, d_stage, nbytes, d_table, d_stats);
    if (rc) return rc;
    if (!h_bed && nbytes) SALSA_FAIL(SALSA_ERR_ARG, "h_bed is NULL");
    if (chunk_bytes < 4 * (size_t)SALSA_MAX_LINE || (chunk_bytes & 15))
        SALSA_FAIL(SALSA_ERR_ARG, "chunk_bytes must be a multiple of 16 and >= %d", 4 * SALSA_MAX_LINE);
    rc = salsa_ctx_streams(ctx);
    if (rc) return rc;
    const char *h = (const char *)h_bed;
    size_t pos = 0;
    int k = 0;
    // the caller's default-stream work on d_table / d_stats (zeroing) must be visible
    CUDA_TRY(cudaDeviceSynchronize());
    while (pos < nbytes) {
        size_t end = pos + chunk_bytes < nbytes ? pos + chunk_bytes : nbytes;
        size_t next = nbytes;
        if (end < nbytes) {
            // last complete line inside [pos, end)
            const char *nl = (const char *)memrchr(h + pos, '\n', end - pos);
            if (!nl || nl == h + pos) SALSA_FAIL(SALSA_ERR_ARG, "BED line longer than the staging chunk near byte %zu", pos);
            const char *pl = (const char *)memrchr(h + pos, '\n', (size_t)(nl - (h + pos)));
            if (!pl) SALSA_FAIL(SALSA_ERR_ARG, "BED line longer than the staging chunk near byte %zu", pos);
            end = (size_t)(nl - h) + 1;
            next = (size_t)(pl - h) + 1;      // the last line is the next chunk's "previous"
        }
        char *dst = (char *)d_stage + (size_t)(k & 1) * chunk_bytes;
        CUDA_TRY(cudaStreamWaitEvent(ctx->s_copy, ctx->ev_comp[k & 1], 0));   // buffer free again
        CUDA_TRY(cudaMemcpyAsync(dst, h + pos, end - pos, cudaMemcpyHostToDevice, ctx->s_copy));
        CUDA_TRY(cudaEventRecord(ctx->ev_copy[k & 1], ctx->s_copy));
        CUDA_TRY(cudaStreamWaitEvent(ctx->s_comp, ctx->ev_copy[k & 1], 0));
        rc = launch_ingest<MODE_DENSE>(ctx, dst, end - pos, pos, d_table, NULL, 0, d_stats, ctx->s_comp);
        if (rc) return rc;
        CUDA_TRY(cudaEventRecord(ctx->ev_comp[k & 1], ctx->s_comp));
        pos = next;
        k++;
    }
    CUDA_TRY(cudaStreamSynchronize(ctx->s_comp));
    return SALSA_OK;
}
