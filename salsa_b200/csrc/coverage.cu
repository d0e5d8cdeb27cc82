// coverage.cu -- Hi-C physical coverage for misassembly detection.
//
// Replaces break_contigs_start.cpp:307-460 (-m yes, once) and
// break_contigs.cpp:199-356 (every iteration):
//   1. BED scan -> per-contig coverage DIFFERENCE arrays (diff[span_start] += 1,
//      diff[span_end + 1] -= 1 for every accepted same-contig read pair) --
//      a Policy of the shared scanner in bed_scan.cuh;
//   2. per segment (a whole contig, or a +-2 Mb window around a scaffold join)
//      an inclusive prefix sum -> coverage, and its exact integer total;
//   3. for the 11 thresholds average/5 .. average/15, Kadane's maximum-sum
//      subarray over delta = (cov < cutoff ? +5 : -5) with the reference's tie
//      rules (first maximum wins, reset only on a strictly negative running
//      sum).  The sequential scan is restated as an associative monoid, so it
//      is a REDUCTION: one pass over the coverage for all 11 thresholds.
//
// Monoid for a segment [a, b] of deltas d (prefix sums p_k = d_a + .. + d_k,
// p_{a-1} = 0):  T = p_b;  (minV, minPos) = first minimum of p over [a-1, b];
// (maxV, maxPos) = first maximum of p over [a, b];  (bestV, bestEnd, bestStart)
// = first maximum over i of m_i = p_i - min_{j in [a-1, i-1]} p_j, with
// bestStart = 1 + first argmin.  Kadane's max_ending_here before the reset is
// exactly m_i, its `s` is 1 + the first argmin (a reset happens exactly when a
// strict new prefix minimum is reached), and `max_so_far < max_ending_here`
// keeps the first maximum.  combine(L, R): cross candidate
// (L.T - L.minV + R.maxV, end = R.maxPos, start = L.minPos + 1); ties go to the
// smaller end, then to the cross candidate (its argmin is the earlier one).
#include <vector>

#include "bed_scan.cuh"

#define SALSA_DATA_MALFORMED_ SALSA_DATA_MALFORMED
#define SALSA_DATA_RANGE_ SALSA_DATA_RANGE

namespace {

using namespace bedscan;

struct CovArgs {
    const uint64_t *covoff;     // [n] element offset of the contig's diff array; ~0 = length < 1 Mb / unknown
    const int64_t *len;         // [n] contig length as `long` (break_contigs*.cpp read lengths with >> long)
    int *diff;
    uint8_t *touched;           // [n] a line landed on this contig (contig2coverage entry exists)
    const uint64_t *join_off;   // joins variant: CSR of kept join positions per contig (sorted)
    const int64_t *join_pos;
    int first_chunk;
};

template <bool JOINS>
struct CovPolicy {
    typedef CovArgs Args;
    static constexpr bool NEED_RFULL = true;
    static constexpr bool LAZY = false;
    static __device__ __forceinline__ bool lazy_candidate(bool, bool) { return true; }
    static constexpr bool AVOID = false;
    // read.substr(0, len-2) == prev_read.substr(0, len-2)   (size_t wrap for len < 2: the whole string)
    static __device__ __forceinline__ int read_cmp_len(int /*rslash*/, int rfull) { return rfull >= 2 ? rfull - 2 : rfull; }
    static __device__ __forceinline__ bool pair(const CovArgs &C, const ScanArgs &A, const uint8_t *s, const Rec &pv,
                                                const Rec &cu, uint64_t line_off, bool prev_is_first)
    {
        if (cu.flags & 16u) { report(A.stats, line_off, SALSA_DATA_LINE_TOO_LONG); return false; }
        // iss >> contig >> start >> end >> read fails -> the reference's loop BREAKS here
        // (break_contigs_start.cpp:321-322, break_contigs.cpp:212-213): reported, the host truncates and reruns
        if ((cu.flags & 7u) < 4u || (cu.flags & 8u)) { report(A.stats, line_off, SALSA_DATA_MALFORMED_); return false; }
        if (cu.id < 0) return false;                               // contig_length[unknown] == 0 < 1000000
        const uint64_t base = C.covoff[cu.id];
        if (base == ~0ull) return false;                           // :325 / :218
        if (!C.touched[cu.id]) C.touched[cu.id] = 1;               // :333-337 / :226-230 (read first: no store storm on a hot line)
        if (C.first_chunk && prev_is_first) return false;          // the first line is never parsed: prev_* are still initial
        if ((pv.flags & 7u) < 4u || (pv.flags & 8u)) return false; // after a malformed line nothing counts (rerun truncates)
        if (pv.id != cu.id) return false;                          // prev_contig == contig
        if (!reads_equal(s, pv, cu)) return false;                // over read_cmp_len() bytes
        const long long ps = rec_start(pv), pe = rec_end(pv), cs = rec_start(cu), ce = rec_end(cu);
        const long long len = C.len[cu.id];
        const long long cutoff = (long long)(int)(len / 10);       // unordered_map<string,int> contig2cutoff
        long long ss, se;
        if (!JOINS) {
            // the 4-way case split of break_contigs_start.cpp:341-370 collapses to this
            if (pe <= ce) { ss = ps; se = ce; } else { ss = cs; se = pe; }
            if (!(ss >= 0 && se < len)) return false;              // :371
            if (se - ss > cutoff) return false;                    // :373
            if (ss >= len + 10 || se + 1 < 0) { report(A.stats, line_off, SALSA_DATA_RANGE_); return false; }   // .at() throws
        } else {
            if (!(pe <= cs)) return false;                         // break_contigs.cpp:235
            ss = ps; se = ce;
            // check_if_inside(contig, prev_start, end + 1): some kept join window [pos-2M, pos+2M] holds the span
            const long long W = 2000000;
            uint64_t lo = C.join_off[cu.id], hi = C.join_off[cu.id + 1];
            const long long want = se + 1 - W;                     // need pos >= want and pos <= ss + W
            while (lo < hi) {
                uint64_t mid = (lo + hi) >> 1;
                if (C.join_pos[mid] < want) lo = mid + 1; else hi = mid;
            }
            if (lo >= C.join_off[cu.id + 1] || C.join_pos[lo] > ss + W) return false;      // :238
            if (se - ss > cutoff) return false;                    // :240
            if (ss < 0 || ss >= len + 10 || se + 1 < 0 || se + 1 >= len + 10) { report(A.stats, line_off, SALSA_DATA_RANGE_); return false; }
        }
        atomicAdd(&C.diff[base + (uint64_t)ss], 1);                // :375-376 / :242-243
        atomicAdd(&C.diff[base + (uint64_t)(se + 1)], -1);
        return true;
    }
};

// ------------------------------------------------------------------ segments
constexpr int CH = 4096;          // elements per chunk
constexpr int CT = 256;           // threads per chunk
constexpr int PER = CH / CT;      // 16
constexpr int NDIV = 11;          // div = 5 .. 15

struct Seg {
    uint64_t src_off, dst_off, seed_idx;   // seed_idx: element of d_diff added before the first element, ~0 = none
    int64_t len, lo, hi;                   // Kadane range [lo, hi)
};

__device__ __forceinline__ int seg_of_chunk(const uint64_t *seg_chunk_off, int nseg, uint64_t g)
{
    int lo = 0, hi = nseg;
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (seg_chunk_off[mid] <= g) lo = mid + 1; else hi = mid;
    }
    return lo - 1;
}

__global__ void __launch_bounds__(CT) chunk_sum_kernel(const int *diff, const Seg *segs, const uint64_t *seg_chunk_off, int nseg,
                                                       unsigned long long *csum)
{
    __shared__ long long red[CT / 32];
    uint64_t g = blockIdx.x;
    int sg = seg_of_chunk(seg_chunk_off, nseg, g);
    Seg S = segs[sg];
    int64_t e0 = (int64_t)(g - seg_chunk_off[sg]) * CH;
    long long acc = 0;
    for (int i = threadIdx.x; i < CH; i += CT) {
        int64_t e = e0 + i;
        if (e < S.len) acc += diff[S.src_off + (uint64_t)e];
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        long long t = 0;
        for (int w = 0; w < CT / 32; w++) t += red[w];
        csum[g] = (unsigned long long)t;
    }
}

// cexcl: exclusive scan of csum over ALL chunks (two's complement), so the prefix
// inside a segment is cexcl[g] - cexcl[first chunk of the segment]
__global__ void __launch_bounds__(CT) chunk_scan_kernel(const int *diff, int *cov, const Seg *segs, const uint64_t *seg_chunk_off,
                                                        int nseg, const unsigned long long *cexcl, unsigned long long *seg_sum)
{
    __shared__ long long wsum[CT / 32];
    __shared__ long long red[CT / 32];
    uint64_t g = blockIdx.x;
    int sg = seg_of_chunk(seg_chunk_off, nseg, g);
    Seg S = segs[sg];
    int64_t e0 = (int64_t)(g - seg_chunk_off[sg]) * CH + (int64_t)threadIdx.x * PER;
    int v[PER];
    long long run = 0;
#pragma unroll
    for (int j = 0; j < PER; j++) {
        int64_t e = e0 + j;
        v[j] = e < S.len ? diff[S.src_off + (uint64_t)e] : 0;
        run += v[j];
    }
    long long inc = run;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        long long o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    long long wb = 0;
    for (int w = 0; w < warp; w++) wb += wsum[w];
    long long base = (long long)(cexcl[g] - cexcl[seg_chunk_off[sg]]) + wb + inc - run;
    if (S.seed_idx != ~0ull) base += diff[S.seed_idx];           // break_contigs.cpp:283-285: tmpcov[start-1] is the RAW diff value
    long long tot = 0;
    int c = (int)base;                                            // the reference accumulates in int
#pragma unroll
    for (int j = 0; j < PER; j++) {
        int64_t e = e0 + j;
        c += v[j];
        if (e < S.len) { cov[S.dst_off + (uint64_t)e] = c; tot += c; }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, d);
    __syncthreads();
    if (lane == 0) red[warp] = tot;
    __syncthreads();
    if (threadIdx.x == 0) {
        long long t = 0;
        for (int w = 0; w < CT / 32; w++) t += red[w];
        atomicAdd(&seg_sum[sg], (unsigned long long)t);
    }
}

struct Summ {
    int T, minV, minPos, maxV, maxPos, bestV, bestEnd, bestStart;   // maxPos < 0: empty
};

__device__ __forceinline__ Summ summ_empty()
{
    Summ s;
    s.T = 0; s.minV = 0; s.minPos = -1; s.maxV = 0; s.maxPos = -1; s.bestV = 0; s.bestEnd = -1; s.bestStart = -1;
    return s;
}

__device__ __forceinline__ Summ summ_combine(const Summ &L, const Summ &R)
{
    if (L.maxPos < 0) return R;
    if (R.maxPos < 0) return L;
    Summ o;
    o.T = L.T + R.T;
    int rmin = L.T + R.minV;
    if (L.minV <= rmin) { o.minV = L.minV; o.minPos = L.minPos; } else { o.minV = rmin; o.minPos = R.minPos; }
    int rmax = L.T + R.maxV;
    if (L.maxV >= rmax) { o.maxV = L.maxV; o.maxPos = L.maxPos; } else { o.maxV = rmax; o.maxPos = R.maxPos; }
    // candidates in tie order: L.best (smallest end), then cross / R.best by end, cross first on equal ends
    int cv = L.T - L.minV + R.maxV, ce = R.maxPos, cs = L.minPos + 1;
    int bv = L.bestV, be = L.bestEnd, bs = L.bestStart;
    int rv = R.bestV, re = R.bestEnd, rs = R.bestStart;
    // pick between cross and R.best first
    if (rv > cv || (rv == cv && re < ce)) { cv = rv; ce = re; cs = rs; }
    if (cv > bv) { bv = cv; be = ce; bs = cs; }
    o.bestV = bv; o.bestEnd = be; o.bestStart = bs;
    return o;
}

__device__ __forceinline__ Summ summ_shfl_down(const Summ &s, int d)
{
    Summ o;
    o.T = __shfl_down_sync(0xffffffffu, s.T, d);
    o.minV = __shfl_down_sync(0xffffffffu, s.minV, d);
    o.minPos = __shfl_down_sync(0xffffffffu, s.minPos, d);
    o.maxV = __shfl_down_sync(0xffffffffu, s.maxV, d);
    o.maxPos = __shfl_down_sync(0xffffffffu, s.maxPos, d);
    o.bestV = __shfl_down_sync(0xffffffffu, s.bestV, d);
    o.bestEnd = __shfl_down_sync(0xffffffffu, s.bestEnd, d);
    o.bestStart = __shfl_down_sync(0xffffffffu, s.bestStart, d);
    return o;
}

// per chunk, per threshold: the monoid summary of the chunk's elements inside [lo, hi)
__global__ void __launch_bounds__(CT) kadane_chunk_kernel(const int *cov, const Seg *segs, const uint64_t *seg_chunk_off, int nseg,
                                                          const unsigned long long *seg_sum, Summ *summ)
{
    __shared__ Summ wres[CT / 32];
    uint64_t g = blockIdx.x;
    int sg = seg_of_chunk(seg_chunk_off, nseg, g);
    Seg S = segs[sg];
    int64_t c0 = (int64_t)(g - seg_chunk_off[sg]) * CH;
    if (c0 + CH <= S.lo || c0 >= S.hi) {                 // chunk entirely outside the Kadane range
        if (threadIdx.x < NDIV) summ[g * NDIV + threadIdx.x] = summ_empty();
        return;
    }
    int64_t e0 = c0 + (int64_t)threadIdx.x * PER;
    int v[PER];
#pragma unroll
    for (int j = 0; j < PER; j++) {
        int64_t e = e0 + j;
        v[j] = (e >= S.lo && e < S.hi) ? cov[S.dst_off + (uint64_t)e] : 0;
    }
    // average = accumulate(cov, 0.0) / size: the sum of ints is exact in a double (< 2^53)
    const double average = (double)(long long)seg_sum[sg] / (double)S.len;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int q = 0; q < NDIV; q++) {
        const double cutoff = average / (double)(q + 5);
        Summ a = summ_empty();
        int T = 0, minV = 0, minPos = (int)(e0 - 1), maxV = 0, maxPos = -1, bestV = 0, bestEnd = -1, bestStart = -1;
        bool any = false;
#pragma unroll
        for (int j = 0; j < PER; j++) {
            int64_t e = e0 + j;
            if (e >= S.lo && e < S.hi) {
                if (!any) { minPos = (int)e - 1; any = true; }
                T += ((double)v[j] < cutoff) ? 1 : -1;           // delta +5 / -5 scaled by 1/5
                int m = T - minV;
                if (bestEnd < 0 || m > bestV) { bestV = m; bestEnd = (int)e; bestStart = minPos + 1; }
                if (maxPos < 0 || T > maxV) { maxV = T; maxPos = (int)e; }
                if (T < minV) { minV = T; minPos = (int)e; }
            }
        }
        if (any) { a.T = T; a.minV = minV; a.minPos = minPos; a.maxV = maxV; a.maxPos = maxPos; a.bestV = bestV; a.bestEnd = bestEnd; a.bestStart = bestStart; }
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            Summ o = summ_shfl_down(a, d);
            if (lane + d < 32) a = summ_combine(a, o);
        }
        if (lane == 0) wres[warp] = a;
        __syncthreads();
        if (threadIdx.x == 0) {
            Summ t = wres[0];
            for (int w = 1; w < CT / 32; w++) t = summ_combine(t, wres[w]);
            summ[g * NDIV + q] = t;
        }
        __syncthreads();
    }
}

// one thread per (segment, threshold): ordered combine of the chunk summaries
__global__ void kadane_final_kernel(const Summ *summ, const uint64_t *seg_chunk_off, int nseg, long long *out_start, long long *out_end)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nseg * NDIV) return;
    int sg = i / NDIV, q = i % NDIV;
    Summ t = summ_empty();
    for (uint64_t g = seg_chunk_off[sg]; g < seg_chunk_off[sg + 1]; g++) t = summ_combine(t, summ[g * NDIV + q]);
    out_start[i] = t.bestStart;
    out_end[i] = t.bestEnd;
}

}   // namespace

// ---------------------------------------------------------------- C ABI
static int cov_scan(salsa_ctx *ctx, bool joins, const void *d_bed, size_t nbytes, uint64_t base_offset, int first_chunk,
                    const uint64_t *d_covoff, const int64_t *d_len, const uint64_t *d_join_off, const int64_t *d_join_pos,
                    int32_t *d_diff, uint8_t *d_touched, uint64_t *d_stats, void *stream)
{
    if (!ctx || (!d_bed && nbytes) || !d_covoff || !d_len || !d_diff || !d_touched || !d_stats || (joins && (!d_join_off || !d_join_pos)))
        SALSA_FAIL(SALSA_ERR_ARG, "salsa_cov_scan: NULL argument");
    if (ctx->n <= 0) SALSA_FAIL(SALSA_ERR_ARG, "salsa_cov_scan: call salsa_names_build first");
    CUDA_TRY(cudaSetDevice(ctx->device));
    CovArgs C{d_covoff, d_len, d_diff, d_touched, d_join_off, d_join_pos, first_chunk};
    if (joins) return bedscan::launch_scan<CovPolicy<true>>(C, ctx->t, d_bed, nbytes, base_offset, d_stats, as_stream(stream));
    return bedscan::launch_scan<CovPolicy<false>>(C, ctx->t, d_bed, nbytes, base_offset, d_stats, as_stream(stream));
}

extern "C" int salsa_cov_scan_start(salsa_ctx *ctx, const void *d_bed, size_t nbytes, uint64_t base_offset, int32_t first_chunk,
                                    const uint64_t *d_covoff, const int64_t *d_len, int32_t *d_diff, uint8_t *d_touched,
                                    uint64_t *d_stats, void *stream)
{
    return cov_scan(ctx, false, d_bed, nbytes, base_offset, first_chunk, d_covoff, d_len, NULL, NULL, d_diff, d_touched, d_stats, stream);
}

extern "C" int salsa_cov_scan_joins(salsa_ctx *ctx, const void *d_bed, size_t nbytes, uint64_t base_offset, int32_t first_chunk,
                                    const uint64_t *d_covoff, const int64_t *d_len, const uint64_t *d_join_off,
                                    const int64_t *d_join_pos, int32_t *d_diff, uint8_t *d_touched, uint64_t *d_stats, void *stream)
{
    return cov_scan(ctx, true, d_bed, nbytes, base_offset, first_chunk, d_covoff, d_len, d_join_off, d_join_pos, d_diff, d_touched,
                    d_stats, stream);
}

static uint64_t seg_chunks(const salsa_cov_seg *s) { return (uint64_t)((s->len + CH - 1) / CH); }

extern "C" size_t salsa_cov_scratch_bytes(const salsa_cov_seg *h_segs, int32_t nseg)
{
    uint64_t nc = 0;
    for (int32_t i = 0; i < nseg; i++) nc += seg_chunks(&h_segs[i]);
    size_t b = 0;
    b += ((size_t)nseg * sizeof(Seg) + 255) & ~(size_t)255;
    b += (((size_t)nseg + 1) * 8 + 255) & ~(size_t)255;          // seg_chunk_off
    b += ((nc + 2) * 8 + 255) & ~(size_t)255;                    // csum / cexcl
    b += ((size_t)nseg * 8 + 255) & ~(size_t)255;                // seg_sum
    b += (nc * NDIV * sizeof(Summ) + 255) & ~(size_t)255;        // summaries
    b += 2 * (((size_t)nseg * NDIV * 8 + 255) & ~(size_t)255);   // results
    return b + 256;
}

extern "C" int salsa_exclusive_scan_u64(uint64_t *d_vals, size_t n, void *stream);

extern "C" int salsa_cov_analyse(const int32_t *d_diff, int32_t *d_cov, const salsa_cov_seg *h_segs, int32_t nseg, void *d_scratch,
                                 size_t scratch_bytes, int64_t *h_sum, int64_t *h_start, int64_t *h_end, void *stream)
{
    if (nseg <= 0) return SALSA_OK;
    if (!d_diff || !d_cov || !h_segs || !d_scratch || !h_sum || !h_start || !h_end) SALSA_FAIL(SALSA_ERR_ARG, "salsa_cov_analyse: NULL argument");
    if (scratch_bytes < salsa_cov_scratch_bytes(h_segs, nseg)) SALSA_FAIL(SALSA_ERR_ARG, "salsa_cov_analyse: scratch too small");
    static_assert(sizeof(Seg) == sizeof(salsa_cov_seg), "segment layout");
    cudaStream_t st = as_stream(stream);
    std::vector<uint64_t> off((size_t)nseg + 1, 0);
    for (int32_t i = 0; i < nseg; i++) {
        if (h_segs[i].len <= 0 || h_segs[i].len > 0x7fffffff) SALSA_FAIL(SALSA_ERR_ARG, "segment %d has length %lld", i, (long long)h_segs[i].len);
        off[i + 1] = off[i] + seg_chunks(&h_segs[i]);
    }
    uint64_t nc = off[nseg];
    if (nc > 0x7fffffffull) SALSA_FAIL(SALSA_ERR_ARG, "too many coverage chunks");
    char *p = (char *)d_scratch;
    auto take = [&](size_t bytes) { char *r = p; p += (bytes + 255) & ~(size_t)255; return r; };
    Seg *d_segs = (Seg *)take((size_t)nseg * sizeof(Seg));
    uint64_t *d_off = (uint64_t *)take(((size_t)nseg + 1) * 8);
    unsigned long long *d_csum = (unsigned long long *)take((nc + 2) * 8);
    unsigned long long *d_segsum = (unsigned long long *)take((size_t)nseg * 8);
    Summ *d_summ = (Summ *)take(nc * NDIV * sizeof(Summ));
    long long *d_rs = (long long *)take((size_t)nseg * NDIV * 8);
    long long *d_re = (long long *)take((size_t)nseg * NDIV * 8);
    CUDA_TRY(cudaMemcpyAsync(d_segs, h_segs, (size_t)nseg * sizeof(Seg), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(d_off, off.data(), ((size_t)nseg + 1) * 8, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemsetAsync(d_segsum, 0, (size_t)nseg * 8, st));
    chunk_sum_kernel<<<(unsigned)nc, CT, 0, st>>>(d_diff, d_segs, d_off, nseg, d_csum);
    int rc = salsa_exclusive_scan_u64((uint64_t *)d_csum, nc, st);
    if (rc) return rc;
    chunk_scan_kernel<<<(unsigned)nc, CT, 0, st>>>(d_diff, d_cov, d_segs, d_off, nseg, d_csum, d_segsum);
    kadane_chunk_kernel<<<(unsigned)nc, CT, 0, st>>>(d_cov, d_segs, d_off, nseg, d_segsum, d_summ);
    kadane_final_kernel<<<(nseg * NDIV + 127) / 128, 128, 0, st>>>(d_summ, d_off, nseg, d_rs, d_re);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(h_sum, d_segsum, (size_t)nseg * 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(h_start, d_rs, (size_t)nseg * NDIV * 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(h_end, d_re, (size_t)nseg * NDIV * 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return SALSA_OK;
}
