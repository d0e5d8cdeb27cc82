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

// The four 4-element groups of a thread: warp w of a chunk owns elements [512 w, 512 w + 512) as four rows of 128, lane l the
// 16 bytes at 4 l of every row -- 16-byte accesses, consecutive across the lanes (round 1 gave each thread 16 consecutive
// elements: every warp load touched 32 different 128-byte lines and the two kernels were bound by the load/store unit).
__device__ __forceinline__ int4 load_quad(const int *p, uint64_t off, int64_t e, int64_t len, bool aligned)
{
    if (aligned && e + 4 <= len) return *reinterpret_cast<const int4 *>(p + off + (uint64_t)e);
    int4 x;
    x.x = e < len ? p[off + (uint64_t)e] : 0;
    x.y = e + 1 < len ? p[off + (uint64_t)e + 1] : 0;
    x.z = e + 2 < len ? p[off + (uint64_t)e + 2] : 0;
    x.w = e + 3 < len ? p[off + (uint64_t)e + 3] : 0;
    return x;
}

__global__ void __launch_bounds__(CT) chunk_sum_kernel(const int *diff, const Seg *segs, const uint64_t *seg_chunk_off, int nseg,
                                                       unsigned long long *csum)
{
    __shared__ long long red[CT / 32];
    uint64_t g = blockIdx.x;
    int sg = seg_of_chunk(seg_chunk_off, nseg, g);
    Seg S = segs[sg];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t e0 = (int64_t)(g - seg_chunk_off[sg]) * CH + warp * (CH / (CT / 32)) + lane * 4;
    const bool al = (S.src_off & 3ull) == 0;
    long long acc = 0;
#pragma unroll
    for (int k = 0; k < PER / 4; k++) {
        const int4 x = load_quad(diff, S.src_off, e0 + k * 128, S.len, al);
        acc += (long long)x.x + x.y + x.z + x.w;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    if (lane == 0) red[warp] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        long long t = 0;
        for (int w = 0; w < CT / 32; w++) t += red[w];
        csum[g] = (unsigned long long)t;
    }
}

// cexcl: exclusive scan of csum over ALL chunks (two's complement), so the prefix
// inside a segment is cexcl[g] - cexcl[first chunk of the segment]
__global__ void __launch_bounds__(CT, 8) chunk_scan_kernel(const int *diff, int *cov, const Seg *segs, const uint64_t *seg_chunk_off,
                                                        int nseg, const unsigned long long *cexcl, unsigned long long *seg_sum)
{
    __shared__ unsigned wsum[CT / 32];
    __shared__ long long red[CT / 32];
    uint64_t g = blockIdx.x;
    int sg = seg_of_chunk(seg_chunk_off, nseg, g);
    Seg S = segs[sg];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t e0 = (int64_t)(g - seg_chunk_off[sg]) * CH + warp * (CH / (CT / 32)) + lane * 4;
    const bool al = ((S.src_off | S.dst_off) & 3ull) == 0;
    int4 v[PER / 4];
    unsigned pre[PER / 4];      // sum of the warp's elements before this thread's group k (mod 2^32: the reference accumulates in int)
    unsigned run = 0;
#pragma unroll
    for (int k = 0; k < PER / 4; k++) {
        v[k] = load_quad(diff, S.src_off, e0 + k * 128, S.len, al);
        const unsigned s = (unsigned)v[k].x + (unsigned)v[k].y + (unsigned)v[k].z + (unsigned)v[k].w;
        unsigned inc = s;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            unsigned o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += o;
        }
        pre[k] = run + inc - s;
        run += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (lane == 0) wsum[warp] = run;
    __syncthreads();
    unsigned wb = 0;
    for (int w = 0; w < warp; w++) wb += wsum[w];
    unsigned base = (unsigned)(cexcl[g] - cexcl[seg_chunk_off[sg]]) + wb;
    if (S.seed_idx != ~0ull) base += (unsigned)diff[S.seed_idx];  // break_contigs.cpp:283-285: tmpcov[start-1] is the RAW diff value
    long long tot = 0;
#pragma unroll
    for (int k = 0; k < PER / 4; k++) {
        const int64_t e = e0 + k * 128;
        int4 c;                                                   // the reference accumulates in int
        c.x = (int)(base + pre[k] + (unsigned)v[k].x);
        c.y = (int)((unsigned)c.x + (unsigned)v[k].y);
        c.z = (int)((unsigned)c.y + (unsigned)v[k].z);
        c.w = (int)((unsigned)c.z + (unsigned)v[k].w);
        if (al && e + 4 <= S.len) {
            *reinterpret_cast<int4 *>(cov + S.dst_off + (uint64_t)e) = c;
            tot += (long long)c.x + c.y + c.z + c.w;
        } else {
            if (e < S.len) { cov[S.dst_off + (uint64_t)e] = c.x; tot += c.x; }
            if (e + 1 < S.len) { cov[S.dst_off + (uint64_t)e + 1] = c.y; tot += c.y; }
            if (e + 2 < S.len) { cov[S.dst_off + (uint64_t)e + 2] = c.z; tot += c.z; }
            if (e + 3 < S.len) { cov[S.dst_off + (uint64_t)e + 3] = c.w; tot += c.w; }
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, d);
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

// per chunk, per threshold: the monoid summary of the chunk's elements inside [lo, hi).
// 64 threads per 4096-element chunk, 64 consecutive elements per thread in registers: the sequential part (about ten
// integer instructions per element and threshold) dominates and the 8-value warp combine is paid once per 64 elements
// (round 1: 256 threads x 16 elements, the combine cost more than the scan; and eleven int -> double conversions per
// element: `(double)cov < cutoff` is `cov < ceil(cutoff)` for an integer cov, computed once per threshold).
constexpr int KT = 64;            // threads per chunk in this kernel
constexpr int KPER = CH / KT;     // 64

__global__ void __launch_bounds__(KT, 10) kadane_chunk_kernel(const int *cov, const Seg *segs, const uint64_t *seg_chunk_off, int nseg,
                                                          const unsigned long long *seg_sum, Summ *summ)
{
    __shared__ Summ wres[KT / 32];
    uint64_t g = blockIdx.x;
    int sg = seg_of_chunk(seg_chunk_off, nseg, g);
    Seg S = segs[sg];
    int64_t c0 = (int64_t)(g - seg_chunk_off[sg]) * CH;
    if (c0 + CH <= S.lo || c0 >= S.hi) {                 // chunk entirely outside the Kadane range
        if (threadIdx.x < NDIV) summ[g * NDIV + threadIdx.x] = summ_empty();
        return;
    }
    const int64_t e0 = c0 + (int64_t)threadIdx.x * KPER;
    // this thread's elements inside [lo, hi): j in [ja, jb)
    const int ja = e0 >= S.lo ? 0 : (S.lo - e0 > KPER ? KPER : (int)(S.lo - e0));
    const int jb = e0 + KPER <= S.hi ? KPER : (S.hi > e0 ? (int)(S.hi - e0) : 0);
    int v[KPER];
    {
        const int4 *src = reinterpret_cast<const int4 *>(cov + S.dst_off + (uint64_t)e0);     // dst_off and e0 are multiples of 4
        const bool whole = e0 + KPER <= S.len && ((S.dst_off & 3ull) == 0);
#pragma unroll
        for (int j = 0; j < KPER; j += 4) {
            if (whole) {
                const int4 x = src[j >> 2];
                v[j] = x.x; v[j + 1] = x.y; v[j + 2] = x.z; v[j + 3] = x.w;
            } else {
#pragma unroll
                for (int t = 0; t < 4; t++) v[j + t] = (e0 + j + t < S.len) ? cov[S.dst_off + (uint64_t)(e0 + j + t)] : 0;
            }
        }
    }
    // average = accumulate(cov, 0.0) / size: the sum of ints is exact in a double (< 2^53)
    const double average = (double)(long long)seg_sum[sg] / (double)S.len;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int q = 0; q < NDIV; q++) {
        const double cutoff = average / (double)(q + 5);
        // (double)c < cutoff  <=>  c < ceil(cutoff) for an int c (never for a NaN cutoff)
        const int ci = !(cutoff == cutoff) ? INT_MIN : (cutoff >= 2147483647.0 ? INT_MAX : (cutoff <= -2147483648.0 ? INT_MIN : (int)ceil(cutoff)));
        const bool all_below = cutoff > 2147483647.0;     // every int is below such a cutoff
        Summ a = summ_empty();
        if (jb > ja) {
            // bestV / maxV start below every reachable value, so the first element always takes them (the reference's
            // "first position wins ties" follows from the strict comparisons)
            int T = 0, minV = 0, minPos = (int)(e0 + ja) - 1, maxV = INT_MIN, maxPos = -1, bestV = INT_MIN, bestEnd = -1, bestStart = -1;
            if (ja == 0 && jb == KPER && (all_below || ci != INT_MIN)) {      // (a cutoff no int is below: the loop further down)
                // value and position in one word, so that every running max / min is ONE integer max / min:
                //   Tk   = T << 14
                //   mink = (minV << 14) + (j_min + 1)                 min: smallest T, then smallest j   (j + 1 = 0: before the thread's first element)
                //   maxk = (maxV << 14) + (127 - j_max)               max: largest T, then smallest j
                //   bstk = (bestV << 14) + ((127 - j_end) << 7) + (126 - j_min)   max: largest sum, then smallest end; the low 7 bits carry
                //          the position of the prefix minimum the sum was taken from (they never decide: (bestV, j_end) is unique)
                // Tk - mink + 127 = ((T - minV) << 14) + (126 - j_min) without a borrow, since 0 <= j_min + 1 <= 64.
                const int cm = all_below ? INT_MAX : ci - 1;              // c < ci  <=>  c <= cm   (ci > INT_MIN: see the branch above)
                constexpr int UP = 1 << 14;
                // Tk holds T << 14 minus (j + 1) * UP, the part every element adds anyway: an element that is NOT below the
                // cutoff then costs one predicated subtraction, and the (j + 1) * UP goes into the constants below
                int Tk = 0, mink = 0, maxk = INT_MIN, bstk = INT_MIN;
#pragma unroll
                for (int j = 0; j < KPER; j++) {
                    if (v[j] > cm) Tk -= 2 * UP;                      // delta +5 / -5 scaled by 1/5
                    bstk = max(bstk, Tk - mink + ((j + 1) * UP + 127 + ((127 - j) << 7)));
                    maxk = max(maxk, Tk + ((j + 1) * UP + (127 - j)));
                    mink = min(mink, Tk + ((j + 1) * UP + (j + 1)));
                }
                Tk += KPER * UP;
                T = Tk >> 14;
                minV = mink >> 14; minPos = (int)e0 + (mink & 0x3FFF) - 1;
                maxV = maxk >> 14; maxPos = (int)e0 + 127 - (maxk & 0x3FFF);
                bestV = bstk >> 14; bestEnd = (int)e0 + 127 - ((bstk & 0x3FFF) >> 7); bestStart = (int)e0 + 126 - (bstk & 127) + 1;
            } else {
                for (int j = ja; j < jb; j++) {
                    const int e = (int)e0 + j;
                    int x = v[0];
#pragma unroll
                    for (int t = 1; t < KPER; t++) x = (t == j) ? v[t] : x;   // registers cannot be indexed
                    T += (all_below || x < ci) ? 1 : -1;
                    const int m = T - minV;
                    if (m > bestV) { bestV = m; bestEnd = e; bestStart = minPos + 1; }
                    if (T > maxV) { maxV = T; maxPos = e; }
                    if (T < minV) { minV = T; minPos = e; }
                }
            }
            a.T = T; a.minV = minV; a.minPos = minPos; a.maxV = maxV; a.maxPos = maxPos; a.bestV = bestV; a.bestEnd = bestEnd; a.bestStart = bestStart;
        }
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            Summ o = summ_shfl_down(a, d);
            if (lane + d < 32) a = summ_combine(a, o);
        }
        if (lane == 0) wres[warp] = a;
        __syncthreads();
        if (threadIdx.x == 0) {
            Summ t = wres[0];
            for (int w = 1; w < KT / 32; w++) t = summ_combine(t, wres[w]);
            summ[g * NDIV + q] = t;
        }
        __syncthreads();
    }
}

// one warp per (segment, threshold): ordered combine of the chunk summaries.  The combine is associative, so lane l
// folds its contiguous share of the segment's chunks and the 32 partial results are combined in lane order
// (round 1: one thread per pair walked all chunks, 1 ms of the analysis for 640,000 chunks).
__global__ void __launch_bounds__(128) kadane_final_kernel(const Summ *summ, const uint64_t *seg_chunk_off, int nseg, long long *out_start,
                                                           long long *out_end)
{
    const int i = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (i >= nseg * NDIV) return;                            // whole warps leave together
    const int sg = i / NDIV, q = i % NDIV;
    const uint64_t g0 = seg_chunk_off[sg], g1 = seg_chunk_off[sg + 1];
    const uint64_t per = (g1 - g0 + 31) / 32;
    uint64_t a = g0 + per * (uint64_t)lane, b = a + per;
    if (a > g1) a = g1;
    if (b > g1) b = g1;
    Summ t = summ_empty();
    for (uint64_t g = a; g < b; g++) t = summ_combine(t, summ[g * NDIV + q]);
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        Summ o = summ_shfl_down(t, d);
        if (lane + d < 32) t = summ_combine(t, o);
    }
    if (lane == 0) {
        out_start[i] = t.bestStart;
        out_end[i] = t.bestEnd;
    }
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
    kadane_chunk_kernel<<<(unsigned)nc, KT, 0, st>>>(d_cov, d_segs, d_off, nseg, d_segsum, d_summ);
    kadane_final_kernel<<<(nseg * NDIV + 3) / 4, 128, 0, st>>>(d_summ, d_off, nseg, d_rs, d_re);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(h_sum, d_segsum, (size_t)nseg * 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(h_start, d_rs, (size_t)nseg * NDIV * 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(h_end, d_re, (size_t)nseg * NDIV * 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return SALSA_OK;
}
