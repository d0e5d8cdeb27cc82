// reproject.cu -- iteration-1 BED -> scaffold coordinates.
//
// Replaces the BED pass of update_bed in layout_unitigs.py:816-886 (after every
// layout) and refactor_breaks.py:133-203 (after every misjoin check): a sliding
// adjacent-line window over alignment_iteration_1.bed that keeps read pairs
// name/1, name/2 whose contigs are both placed and rewrites them as
//     scaffold \t new_start \t new_end \t read \n .
// Three steps, all on the device:
//   1. the shared BED scanner (bed_scan.cuh) with a policy that records the file
//      offset of every line that closes a kept pair (layout variant) or that opens
//      a new same-contig run (refactor variant: its `continue` on equal contigs
//      leaves prev_line at the first line of the previous run, :150-151, so the
//      evaluated pairs are exactly the consecutive run starts);
//   2. a radix sort of those offsets (file order = output order);
//   3. one thread per candidate, one block per 256 consecutive candidates: parse
//      the two lines, map the coordinates, and either return the output length
//      (pass 1: per candidate and per block) or, with the block offsets scanned,
//      write the two lines at their place (pass 2).
// HBM traffic: the BED once in step 1; steps 2-3 touch only candidate lines.
#include <cub/cub.cuh>

#include "bed_scan.cuh"

namespace {

using namespace bedscan;

struct CandArgs {
    const int32_t *scaf_of;          // contig id -> scaffold index, -1 = not placed
    uint64_t *cand;
    uint64_t cap;
    unsigned long long *count;
};

__device__ __forceinline__ void append(const CandArgs &C, uint64_t line_off)
{
    unsigned m = __activemask();
    int lane = threadIdx.x & 31;
    int leader = __ffs(m) - 1;
    unsigned long long base = 0;
    if (lane == leader) base = atomicAdd(C.count, (unsigned long long)__popc(m));
    base = __shfl_sync(m, base, leader);
    unsigned long long idx = base + __popc(m & ((1u << lane) - 1u));
    if (idx < C.cap) C.cand[idx] = line_off;
}

// token.split('/')[1] == ch
__device__ __forceinline__ bool mate_is(const uint8_t *s, const Rec &r, uint32_t ch)
{
    int roff = (int)(r.read & 0xFFFF), rs = (int)(r.read >> 16), rf = (int)((r.flags >> 8) & 0xFFFF);
    return rs + 1 < rf && s[roff + rs + 1] == ch && (rs + 2 == rf || s[roff + rs + 2] == '/');
}

template <bool REFAC>
struct CandPolicy {
    typedef CandArgs Args;
    static constexpr bool AVOID = false;
    static constexpr bool NEED_RFULL = true;
    static constexpr bool LAZY = false;
    static __device__ __forceinline__ bool lazy_candidate(bool, bool) { return true; }
    static __device__ __forceinline__ int read_cmp_len(int rslash, int /*rfull*/) { return rslash; }

    static __device__ __forceinline__ bool pair(const CandArgs &C, const ScanArgs &A, const uint8_t *s, const Rec &pv,
                                                const Rec &cu, uint64_t line_off, bool /*prev_is_first*/)
    {
        uint32_t pnf = pv.flags & 7u, cnf = cu.flags & 7u;
        if ((pv.flags | cu.flags) & 16u) { report(A.stats, line_off, SALSA_DATA_LINE_TOO_LONG); return false; }
        if (REFAC) {
            if (pnf < 1 || cnf < 1) { report(A.stats, line_off, SALSA_DATA_INDEX_ERROR); return false; }   // refactor_breaks.py:146-147
            if (same_contig(s, pv, cu)) return false;                                                      // :150-151
            append(C, line_off);
            return true;
        }
        if (pnf < 4 || cnf < 4) { report(A.stats, line_off, SALSA_DATA_INDEX_ERROR); return false; }       // layout_unitigs.py:836-841
        uint32_t prs = pv.read >> 16, prf = (pv.flags >> 8) & 0xFFFFu, crs = cu.read >> 16, crf = (cu.flags >> 8) & 0xFFFFu;
        if (prs == prf || crs == crf) { report(A.stats, line_off, SALSA_DATA_INDEX_ERROR); return false; } // split('/')[1], :842-843
        if (pv.id < 0 || cu.id < 0) return false;                                                          // :844
        if (C.scaf_of[pv.id] < 0 || C.scaf_of[cu.id] < 0) return false;
        if (!reads_equal(s, pv, cu)) return false;                                                         // :847
        if (!mate_is(s, pv, '1') || !mate_is(s, cu, '2')) return false;
        if ((pv.flags | cu.flags) & 8u) { report(A.stats, line_off, SALSA_DATA_VALUE_ERROR); return false; }  // int(), :850-851
        append(C, line_off);
        return true;
    }
};

// ---- step 3: per-candidate evaluation straight from global memory ----
struct EmitArgs {
    const uint8_t *bed;
    uint64_t nbytes;
    const uint64_t *cand;
    uint64_t ncand;
    DevTable t;
    const int32_t *scaf_of;
    const int64_t *off, *len;
    const uint8_t *rev;
    const char *sblob;
    const uint32_t *soff;
    uint32_t *len32;
    uint64_t *blk;
    uint8_t *out;
    uint64_t out_cap;
    uint64_t *stats;
};

struct GLine {
    int nf, id, nlen, tlen, s1, s2;      // s1 / s2: first / second '/' inside the read token (tlen if none)
    bool bad;
    long long a, b;
    uint64_t tok;
};

__device__ __forceinline__ uint32_t gbyte(const EmitArgs &E, uint64_t q) { return q < E.nbytes ? E.bed[q] : (uint32_t)'\n'; }

// str.split() of one line starting at p, fields as update_bed uses them
__device__ GLine gparse(const EmitArgs &E, uint64_t p)
{
    GLine r;
    r.nf = 0; r.id = -1; r.nlen = 0; r.tlen = 0; r.s1 = r.s2 = 0; r.bad = false; r.a = r.b = 0; r.tok = 0;
    uint32_t c = gbyte(E, p);
    while (c != '\n' && is_ws(c)) c = gbyte(E, ++p);
    if (c == '\n') return r;
    uint64_t noff = p;
    uint32_t w[4] = {0, 0, 0, 0};
    int nlen = 0;
    while (!is_ws(c)) {
        if (nlen < 16) w[nlen >> 2] |= c << ((nlen & 3) * 8);
        nlen++;
        c = gbyte(E, ++p);
    }
    r.nlen = nlen;
    r.id = lookup_words(E.t, E.bed + noff, w[0], w[1], w[2], w[3], 0, nlen);
    r.nf = 1;
    long long val[2] = {0, 0};
    for (int f = 0; f < 2; f++) {
        while (c != '\n' && is_ws(c)) c = gbyte(E, ++p);
        if (c == '\n') { r.a = val[0]; r.b = val[1]; return r; }
        r.nf++;
        bool neg = (c == '-');
        if (c == '-' || c == '+') c = gbyte(E, ++p);
        unsigned long long v = 0;
        int nd = 0;
        uint32_t d = c - '0';
        while (d <= 9u) { v = v * 10ull + d; nd++; c = gbyte(E, ++p); d = c - '0'; }
        if (nd == 0 || nd > 18 || !is_ws(c)) r.bad = true;
        while (!is_ws(c)) c = gbyte(E, ++p);
        val[f] = neg ? -(long long)v : (long long)v;
    }
    r.a = val[0]; r.b = val[1];
    while (c != '\n' && is_ws(c)) c = gbyte(E, ++p);
    if (c == '\n') return r;
    r.nf = 4;
    r.tok = p;
    int k = 0, s1 = -1, s2 = -1;
    while (!is_ws(c)) {
        if (c == '/') { if (s1 < 0) s1 = k; else if (s2 < 0) s2 = k; }
        k++;
        c = gbyte(E, ++p);
    }
    r.tlen = k;
    r.s1 = s1 < 0 ? k : s1;
    r.s2 = s2 < 0 ? k : s2;
    return r;
}

__device__ __forceinline__ int ndigits(long long v)
{
    unsigned long long u = v < 0 ? 0ull - (unsigned long long)v : (unsigned long long)v;
    int n = v < 0 ? 2 : 1;
    if (u >> 32) {
        while (u >= 10ull) { u /= 10ull; n++; }
        return n;
    }
    uint32_t w = (uint32_t)u;
    return n + (w >= 10u) + (w >= 100u) + (w >= 1000u) + (w >= 10000u) + (w >= 100000u) + (w >= 1000000u) + (w >= 10000000u) +
           (w >= 100000000u) + (w >= 1000000000u);
}

// Output bytes of one thread: 4-byte aligned words once the address is aligned (the words before and after
// are shared with the neighbouring candidates' output, so the head and the tail go out byte by byte).
struct Out {
    uint8_t *p;
    uint32_t acc;
    bool al;
    __device__ __forceinline__ void init(uint8_t *o) { p = o; acc = 0; al = false; }
    __device__ __forceinline__ void put(uint32_t b)
    {
        uint32_t k = (uint32_t)(uintptr_t)p & 3u;
        if (k == 0) al = true;
        if (!al) *p = (uint8_t)b;
        else {
            acc |= b << (8 * k);
            if (k == 3) { *reinterpret_cast<uint32_t *>(p - 3) = acc; acc = 0; }
        }
        p++;
    }
    __device__ __forceinline__ void flush()
    {
        uint32_t k = (uint32_t)(uintptr_t)p & 3u;
        if (al) for (uint32_t j = 0; j < k; j++) p[(int)j - (int)k] = (uint8_t)(acc >> (8 * j));
    }
};

__device__ __forceinline__ void put_int(Out &o, long long v)
{
    unsigned long long u = v < 0 ? 0ull - (unsigned long long)v : (unsigned long long)v;
    if (v < 0) o.put('-');
    if (u >> 32) {                                       // rare: beyond 4.29e9
        char buf[20];
        int n = 0;
        do { buf[n++] = (char)('0' + (int)(u % 10ull)); u /= 10ull; } while (u);
        while (n) o.put((uint32_t)buf[--n]);
        return;
    }
    uint32_t w = (uint32_t)u;
    unsigned long long bcd = 0;                          // digits as nibbles, most significant in the lowest nibble
    int n = 0;
    do { uint32_t q = w / 10u; bcd = (bcd << 4) | (w - q * 10u); w = q; n++; } while (w);
    for (; n; n--) { o.put('0' + (uint32_t)(bcd & 15ull)); bcd >>= 4; }
}

__device__ __forceinline__ void mapped(const EmitArgs &E, const GLine &g, long long &a, long long &b)
{
    long long off = E.off[g.id], len = E.len[g.id];
    if (E.rev[g.id]) { a = off + len - g.b; b = off + len - g.a; }       // layout_unitigs.py:858-863
    else { a = off + g.a; b = off + g.b; }
}

__device__ __forceinline__ void put_line(const EmitArgs &E, Out &o, const GLine &g, int sc, long long a, long long b)
{
    uint32_t s0 = E.soff[sc], s1 = E.soff[sc + 1];
    for (uint32_t i = s0; i < s1; i++) o.put((uint8_t)E.sblob[i]);
    o.put('\t');
    put_int(o, a);
    o.put('\t');
    put_int(o, b);
    o.put('\t');
    for (int i = 0; i < g.tlen; i++) o.put(E.bed[g.tok + i]);
    o.put('\n');
}

constexpr int EB = 256;                 // candidates per block

// One thread per candidate, one block per EB consecutive candidates (= a contiguous piece of the output).
// WRITE = false: byte count per candidate -> E.len32, per block -> E.blk.  WRITE = true: E.blk holds the
// scanned block offsets; a block-wide scan of the counts places every candidate's two lines.
template <bool REFAC, bool WRITE>
__global__ void __launch_bounds__(EB) reproj_emit_kernel(EmitArgs E)
{
    __shared__ uint32_t wsum[EB / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint64_t k = (uint64_t)blockIdx.x * EB + tid;
    uint32_t n = 0;
    GLine p, c;
    int sp = -1, sc = -1;
    long long pa = 0, pb = 0, ca = 0, cb = 0;
    if (k < E.ncand) {
        uint64_t cur = E.cand[k], prev;
        if (REFAC) {
            prev = k ? E.cand[k - 1] : 0;                   // the file's first line opens the first run
        } else {
            prev = cur >= 2 ? cur - 2 : 0;                  // cur - 1 is the previous line's '\n'
            while (prev > 0 && E.bed[prev] != '\n') prev--;
            if (E.bed[prev] == '\n' && prev + 1 < cur) prev++;
        }
        p = gparse(E, prev);
        c = gparse(E, cur);
        if (p.nf < 4 || c.nf < 4 || p.s1 == p.tlen || c.s1 == c.tlen) {
            report(E.stats, cur, SALSA_DATA_INDEX_ERROR);   // pa[3] / split('/')[1]
        } else if (p.id >= 0 && c.id >= 0) {
            sp = E.scaf_of[p.id];
            sc = E.scaf_of[c.id];
            if (sp >= 0 && sc >= 0 && p.s1 == c.s1 && p.s2 == p.s1 + 2 && c.s2 == c.s1 + 2 &&
                E.bed[p.tok + p.s1 + 1] == '1' && E.bed[c.tok + c.s1 + 1] == '2' && (!REFAC || sp != sc)) {
                bool keep = true;
                for (int i = 0; i < c.s1; i++)
                    if (E.bed[p.tok + i] != E.bed[c.tok + i]) { keep = false; break; }
                if (keep && (p.bad || c.bad)) { report(E.stats, cur, SALSA_DATA_VALUE_ERROR); keep = false; }
                if (keep) {
                    mapped(E, p, pa, pb);
                    mapped(E, c, ca, cb);
                    n = (E.soff[sp + 1] - E.soff[sp]) + (E.soff[sc + 1] - E.soff[sc]) + 8u +
                        (uint32_t)(ndigits(pa) + ndigits(pb) + ndigits(ca) + ndigits(cb) + p.tlen + c.tlen);
                }
            }
        }
    }
    // block-wide inclusive scan of n
    uint32_t x = n;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t o = __shfl_up_sync(0xffffffffu, x, d);
        if (lane >= d) x += o;
    }
    if (lane == 31) wsum[warp] = x;
    __syncthreads();
    uint32_t before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < EB / 32; w++) {
        uint32_t v = wsum[w];
        if (w < warp) before += v;
        total += v;
    }
    if (!WRITE) {
        if (k < E.ncand) E.len32[k] = n;
        if (tid == 0) E.blk[blockIdx.x] = total;
        return;
    }
    const uint32_t lo = before + x - n;
    const uint64_t gbase = E.blk[blockIdx.x];
    if (gbase + total > E.out_cap) return;                  // the caller sized the output from the same counts
    if (n) {
        // direct byte stores: measured faster than assembling the block's piece in shared memory first, whose
        // 40 KB per block shrink the L1 that the per-thread byte loads of the two source lines live in
        Out o;
        o.init(E.out + gbase + lo);
        put_line(E, o, p, sp, pa, pb);
        put_line(E, o, c, sc, ca, cb);
        o.flush();
    }
}

}   // namespace

extern "C" int salsa_reproject_candidates(salsa_ctx *ctx, const void *d_bed, size_t nbytes, int32_t refactor,
                                          const int32_t *d_scaffold_of, uint64_t *d_cand, size_t capacity,
                                          uint64_t *d_count, uint64_t *d_stats, void *stream)
{
    if (!ctx || (!d_bed && nbytes) || !d_scaffold_of || (!d_cand && capacity) || !d_count || !d_stats)
        SALSA_FAIL(SALSA_ERR_ARG, "salsa_reproject_candidates: NULL argument");
    if (ctx->n <= 0) SALSA_FAIL(SALSA_ERR_ARG, "salsa_reproject_candidates: call salsa_names_build first");
    CUDA_TRY(cudaSetDevice(ctx->device));
    CUDA_TRY(cudaMemsetAsync(d_count, 0, sizeof(uint64_t), as_stream(stream)));
    CandArgs C{d_scaffold_of, d_cand, capacity, (unsigned long long *)d_count};
    DevTable t = ctx->t;
    t.navoid = 0;
    if (refactor) return bedscan::launch_scan<CandPolicy<true>>(C, t, d_bed, nbytes, 0, d_stats, as_stream(stream));
    return bedscan::launch_scan<CandPolicy<false>>(C, t, d_bed, nbytes, 0, d_stats, as_stream(stream));
}

extern "C" int salsa_sort_u64(uint64_t *d_in, uint64_t *d_out, size_t n, uint64_t max_key, void *d_tmp, size_t *tmp_bytes, void *stream)
{
    if (!tmp_bytes) SALSA_FAIL(SALSA_ERR_ARG, "tmp_bytes is NULL");
    if (n > 0x7fffffffull) SALSA_FAIL(SALSA_ERR_ARG, "too many keys to sort");
    // only the bits a key can have are sorted: line offsets of a 10 GB BED need 34, which is 5 radix passes instead of 8
    int bits = 64;
    if (max_key) { bits = 1; while (bits < 64 && (max_key >> bits)) bits++; }
    size_t need = 0;
    CUDA_TRY(cub::DeviceRadixSort::SortKeys(NULL, need, d_in, d_out, (int)n, 0, bits, as_stream(stream)));
    if (!d_tmp) { *tmp_bytes = need; return SALSA_OK; }
    if (*tmp_bytes < need) SALSA_FAIL(SALSA_ERR_ARG, "sort scratch too small: need %zu", need);
    CUDA_TRY(cub::DeviceRadixSort::SortKeys(d_tmp, need, d_in, d_out, (int)n, 0, bits, as_stream(stream)));
    return SALSA_OK;
}

extern "C" size_t salsa_reproject_blocks(size_t ncand) { return (ncand + EB - 1) / EB; }

extern "C" int salsa_reproject_emit(salsa_ctx *ctx, const void *d_bed, size_t nbytes, int32_t refactor,
                                    const uint64_t *d_cand, size_t ncand, const int32_t *d_scaffold_of,
                                    const int64_t *d_offset, const int64_t *d_length, const uint8_t *d_reversed,
                                    const char *d_scaf_names, const uint32_t *d_scaf_name_off, uint32_t *d_len,
                                    uint64_t *d_blk, void *d_out, size_t out_capacity, uint64_t *d_stats, void *stream)
{
    if (!ctx || (!d_bed && nbytes) || (!d_cand && ncand) || !d_scaffold_of || !d_offset || !d_length || !d_reversed ||
        !d_scaf_names || !d_scaf_name_off || !d_len || !d_blk || !d_stats)
        SALSA_FAIL(SALSA_ERR_ARG, "salsa_reproject_emit: NULL argument");
    if (ctx->n <= 0) SALSA_FAIL(SALSA_ERR_ARG, "salsa_reproject_emit: call salsa_names_build first");
    cudaStream_t st = as_stream(stream);
    uint64_t nb = (ncand + EB - 1) / EB;
    if (nb > 0x7fffffffull) SALSA_FAIL(SALSA_ERR_ARG, "salsa_reproject_emit: too many candidates");
    CUDA_TRY(cudaSetDevice(ctx->device));
    if (!d_out) CUDA_TRY(cudaMemsetAsync(d_blk + nb, 0, sizeof(uint64_t), st));
    if (ncand == 0) return SALSA_OK;
    EmitArgs E{(const uint8_t *)d_bed, nbytes, d_cand, ncand, ctx->t, d_scaffold_of, d_offset, d_length, d_reversed,
               d_scaf_names, d_scaf_name_off, d_len, d_blk, (uint8_t *)d_out, out_capacity, d_stats};
    if (refactor) {
        if (d_out) reproj_emit_kernel<true, true><<<(unsigned)nb, EB, 0, st>>>(E);
        else reproj_emit_kernel<true, false><<<(unsigned)nb, EB, 0, st>>>(E);
    } else {
        if (d_out) reproj_emit_kernel<false, true><<<(unsigned)nb, EB, 0, st>>>(E);
        else reproj_emit_kernel<false, false><<<(unsigned)nb, EB, 0, st>>>(E);
    }
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}
