// links_score.cu -- enzyme-normalised link scores and best-buddy ratios.
//
// Replaces, for every link of the table:
//   make_links.py:104-137,165-168   norm = RE[first]*0.5 + RE[second]*0.5, score = count/norm
//   make_links.py:169 -> fast_scaled_scores.py:48,71   the score crosses a text file as
//        Python-2 str(float) = '%.12g', so the value the next program sees is the
//        double nearest to the 12-significant-digit decimal of the score
//   fast_scaled_scores.py:47-57     node2weight[end] = max weight over incident edges
//   fast_scaled_scores.py:13-33,79-86  bestAlt / ratio (best-buddy second-best ratio)
//
// The decimal round trip is done exactly, with 320-bit integers on the device:
// g12(x) yields the 12-digit integer D and decimal exponent k of '%.12g' % x
// (round-half-even on the exact binary value, like glibc/CPython) and the
// correctly rounded double of D*10^(k-11).  The host only assembles text from
// (D, k); it never converts a float.
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "common.cuh"

namespace {

constexpr int NL = 10;   // 32-bit limbs: 320 bits (largest intermediate < 2^250)

// `digits` values that are not digits (include/salsa_b200.h: salsa_g12)
#define SALSA_G12_RANGE INT64_MIN
#define SALSA_G12_INF (INT64_MIN + 1)
#define SALSA_G12_NINF (INT64_MIN + 2)
#define SALSA_G12_NAN (INT64_MIN + 3)

struct Big {
    uint32_t w[NL];
};

__device__ __forceinline__ void big_set(Big &a, uint64_t v)
{
#pragma unroll
    for (int i = 0; i < NL; i++) a.w[i] = 0;
    a.w[0] = (uint32_t)v;
    a.w[1] = (uint32_t)(v >> 32);
}

__device__ __forceinline__ void big_mul_small(Big &a, uint32_t m)
{
    uint64_t c = 0;
#pragma unroll
    for (int i = 0; i < NL; i++) {
        uint64_t t = (uint64_t)a.w[i] * m + c;
        a.w[i] = (uint32_t)t;
        c = t >> 32;
    }
}

__device__ void big_mul_pow10(Big &a, int p)
{
    while (p >= 9) { big_mul_small(a, 1000000000u); p -= 9; }
    uint32_t m = 1;
    while (p-- > 0) m *= 10u;
    if (m > 1) big_mul_small(a, m);
}

__device__ void big_shl(Big &a, int bits)
{
    int ws = bits >> 5, bs = bits & 31;
    Big r;
#pragma unroll
    for (int i = 0; i < NL; i++) {
        uint32_t lo = (i - ws >= 0) ? a.w[(i - ws) < 0 ? 0 : (i - ws)] : 0u;
        uint32_t lo1 = (i - ws - 1 >= 0) ? a.w[(i - ws - 1) < 0 ? 0 : (i - ws - 1)] : 0u;
        r.w[i] = bs ? ((lo << bs) | (lo1 >> (32 - bs))) : lo;
    }
    a = r;
}

__device__ __forceinline__ void big_shr1(Big &a)
{
#pragma unroll
    for (int i = 0; i < NL - 1; i++) a.w[i] = (a.w[i] >> 1) | (a.w[i + 1] << 31);
    a.w[NL - 1] >>= 1;
}

__device__ __forceinline__ int big_cmp(const Big &a, const Big &b)
{
    for (int i = NL - 1; i >= 0; i--) {
        if (a.w[i] != b.w[i]) return a.w[i] > b.w[i] ? 1 : -1;
    }
    return 0;
}

__device__ __forceinline__ void big_sub(Big &a, const Big &b)
{
    uint32_t borrow = 0;
#pragma unroll
    for (int i = 0; i < NL; i++) {
        uint64_t t = (uint64_t)a.w[i] - b.w[i] - borrow;
        a.w[i] = (uint32_t)t;
        borrow = (uint32_t)(t >> 63);
    }
}

__device__ __forceinline__ bool big_is_zero(const Big &a)
{
    uint32_t o = 0;
#pragma unroll
    for (int i = 0; i < NL; i++) o |= a.w[i];
    return o == 0;
}

__device__ int big_bitlen(const Big &a)
{
    for (int i = NL - 1; i >= 0; i--)
        if (a.w[i]) return i * 32 + (32 - __clz(a.w[i]));
    return 0;
}

// floor(N / M) for a quotient known to be < 2^qbits; N becomes the remainder.
__device__ uint64_t big_divq(Big &N, const Big &M, int qbits)
{
    Big T = M;
    big_shl(T, qbits - 1);
    uint64_t q = 0;
    for (int i = qbits - 1; i >= 0; i--) {
        if (big_cmp(N, T) >= 0) { big_sub(N, T); q |= 1ull << i; }
        big_shr1(T);
    }
    return q;
}

// round-half-even of N/M given the floor q and the remainder R (in place of N)
__device__ __forceinline__ uint64_t round_half_even(uint64_t q, Big &R, const Big &M)
{
    big_shl(R, 1);
    int c = big_cmp(R, M);
    if (c > 0 || (c == 0 && (q & 1))) q++;
    return q;
}

// x > 0, finite, normal.  Returns D (10^11 <= D < 10^12), *k10 = decimal exponent of
// the leading digit, *w12 = double nearest to D * 10^(k10 - 11).
__device__ uint64_t g12(double x, int *k10, double *w12)
{
    int e2;
    double fr = frexp(x, &e2);                 // x = fr * 2^e2, 0.5 <= fr < 1
    uint64_t m = (uint64_t)ldexp(fr, 53);      // 53-bit integer
    int e = e2 - 53;                           // x = m * 2^e
    int k = (int)floor(log10(x));
    uint64_t D = 0;
    Big N, M;
    for (int iter = 0; iter < 4; iter++) {
        int q = k - 11;
        big_set(N, m);
        big_set(M, 1);
        if (e >= 0) big_shl(N, e); else big_shl(M, -e);
        if (q >= 0) big_mul_pow10(M, q); else big_mul_pow10(N, -q);
        D = big_divq(N, M, 45);
        if (D >= 1000000000000ull) { k++; continue; }
        if (D < 100000000000ull) { k--; continue; }
        break;
    }
    D = round_half_even(D, N, M);
    if (D == 1000000000000ull) { D = 100000000000ull; k++; }
    *k10 = k;
    // ---- D * 10^q -> nearest double --------------------------------------
    int q = k - 11;
    big_set(N, D);
    big_set(M, 1);
    if (q >= 0) big_mul_pow10(N, q); else big_mul_pow10(M, -q);
    int s = 53 - (big_bitlen(N) - big_bitlen(M));    // quotient of (N << s) / M has 53 or 54 bits
    uint64_t Q = 0;
    Big N2;
    for (int iter = 0; iter < 2; iter++) {
        N2 = N;
        Big M2 = M;
        if (s >= 0) big_shl(N2, s); else big_shl(M2, -s);
        Q = big_divq(N2, M2, 55);
        if (Q >= (1ull << 53)) { s--; continue; }
        Q = round_half_even(Q, N2, M2);
        break;
    }
    *w12 = ldexp((double)Q, -s);
    return D;
}

// 10^p, p = 0..19 (each also exactly representable as a double: 5^19 < 2^53)
__constant__ uint64_t P10[20] = {1ull, 10ull, 100ull, 1000ull, 10000ull, 100000ull, 1000000ull, 10000000ull, 100000000ull,
                                 1000000000ull, 10000000000ull, 100000000000ull, 1000000000000ull, 10000000000000ull,
                                 100000000000000ull, 1000000000000000ull, 10000000000000000ull, 100000000000000000ull,
                                 1000000000000000000ull, 10000000000000000000ull};

// The same result as g12() for 1e-8 <= x < 1e12 (every score and ratio the pipeline produces), in a handful of
// instructions: with x = m * 2^-s (m < 2^53) and p = 11 - k in 0..19, the integer m * 10^p fits 128 bits, so
// D = round-half-even(m * 10^p / 2^s) is a shift with an exact look at the bits shifted out; and the double nearest
// to D / 10^p is ONE IEEE division, because D < 2^53 and 10^p (p <= 22) are both exact doubles (Clinger's fast path).
__device__ __forceinline__ bool g12_fast(double x, uint64_t *Dout, int *k10, double *w12)
{
    if (!(x >= 1e-8 && x < 1e12)) return false;
    const uint64_t bits = (uint64_t)__double_as_longlong(x);
    const int be = (int)(bits >> 52);                              // x >= 1e-8: normal
    const uint64_t m = (bits & 0xFFFFFFFFFFFFFull) | (1ull << 52);
    const int s = 1075 - be;                                       // x = m * 2^-s, 13 <= s <= 80
    int k = (int)floor(log10(x));
    uint64_t D = 0;
    for (int iter = 0; iter < 3; iter++) {
        const int p = 11 - k;
        if (p < 0 || p > 19) return false;
        const uint64_t T = P10[p];
        const uint64_t lo = m * T, hi = __umul64hi(m, T);          // m * 10^p, 128 bits
        uint64_t half, sticky;
        if (s < 64) {
            D = (lo >> s) | (hi << (64 - s));
            if (hi >> s) return false;                             // does not happen (D would exceed 64 bits)
            half = (lo >> (s - 1)) & 1ull;
            sticky = lo & ((1ull << (s - 1)) - 1ull);
        } else if (s == 64) {
            D = hi;
            half = lo >> 63;
            sticky = lo & 0x7FFFFFFFFFFFFFFFull;
        } else {
            D = hi >> (s - 64);
            half = (hi >> (s - 65)) & 1ull;
            sticky = (hi & ((1ull << (s - 65)) - 1ull)) | lo;
        }
        if (D >= 1000000000000ull) { k++; continue; }              // log10 was one off
        if (D < 100000000000ull) { k--; continue; }
        if (half && (sticky || (D & 1ull))) D++;                   // round half even on the exact value
        if (D == 1000000000000ull) { D = 100000000000ull; k++; }
        *Dout = D;
        *k10 = k;
        const int q = k - 11;                                      // D * 10^q -> nearest double
        *w12 = q <= 0 ? __ddiv_rn((double)D, (double)P10[-q]) : __dmul_rn((double)D, (double)P10[q]);
        return q >= -19 && q <= 0;
    }
    return false;
}

__device__ __forceinline__ void g12_any(double x, int64_t *digits, int32_t *exp10, double *w12)
{
    if (x == 0.0) {                   // prints '0.0'
        *digits = 0; *exp10 = 0; *w12 = x;
        return;
    }
    if (!isfinite(x)) {               // cannot arise from count/norm; an arbitrary input file can hold them: 'inf', '-inf', 'nan'
        *digits = isnan(x) ? SALSA_G12_NAN : (x > 0 ? SALSA_G12_INF : SALSA_G12_NINF); *exp10 = 0; *w12 = x;
        return;
    }
    double a = fabs(x), w;
    if (a < 1e-30 || a > 1e30) {      // outside the 320-bit working range: flagged, the formatters refuse it
        *digits = SALSA_G12_RANGE; *exp10 = 0; *w12 = x;
        return;
    }
    int k;
    uint64_t D;
    if (!g12_fast(a, &D, &k, &w)) D = g12(a, &k, &w);               // the exact 320-bit route covers the rest of the range
    *digits = x < 0 ? -(int64_t)D : (int64_t)D;
    *exp10 = k;
    *w12 = x < 0 ? -w : w;
}

__global__ void g12_kernel(const double *x, size_t n, int64_t *digits, int32_t *exp10, double *w12)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    g12_any(x[i], &digits[i], &exp10[i], &w12[i]);
}

// make_links.py:104-137,165-168
__global__ void score_kernel(const uint64_t *keys, const uint64_t *counts, size_t n, uint32_t ncontig,
                             const int64_t *re_left, const int64_t *re_right, double *score, uint8_t *is_int0,
                             int64_t *digits, int32_t *exp10, double *w12, int32_t *u, int32_t *v)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t key = keys[i];
    uint32_t ends = (uint32_t)(key & 3u);
    uint64_t pr = key >> 2;
    uint32_t first = (uint32_t)(pr / ncontig), second = (uint32_t)(pr % ncontig);
    double x = (double)((ends & 2u) ? re_right[first] : re_left[first]);
    double y = (double)((ends & 1u) ? re_right[second] : re_left[second]);
    // r1 = (len/2)/len == 0.5 exactly for every non-zero finite length
    double norm = __dadd_rn(__dmul_rn(x, 0.5), __dmul_rn(y, 0.5));
    double s;
    uint8_t int0 = 0;
    if (norm == 0.0) { s = 0.0; int0 = 1; }                       // the int literal 0, :165-166
    else s = __ddiv_rn(__dmul_rn((double)counts[i], 1.0), norm);  // :168
    score[i] = s;
    is_int0[i] = int0;
    g12_any(s, &digits[i], &exp10[i], &w12[i]);
    u[i] = (int32_t)(first * 2u + ((ends >> 1) & 1u));
    v[i] = (int32_t)(second * 2u + (ends & 1u));
}

// ---- best-buddy: per-node top-2 with multiplicity ------------------------
// weights are >= 0, so the IEEE bit pattern orders like the value.
// The edge list is key-sorted, so the 32 edges of a warp share two or three `u` nodes (one contig, two ends) while their
// `v` nodes are spread out: the u side is reduced inside the warp first (one atomic per distinct node and warp instead of
// 32 atomics on the same address), the v side goes straight to the atomics.
__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long x)
{
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        unsigned long long o = __shfl_xor_sync(0xffffffffu, x, d);
        x = o > x ? o : x;
    }
    return x;
}

__global__ void node_max_kernel(const int32_t *u, const int32_t *v, const double *w, size_t n, uint32_t nnodes, unsigned long long *nmax)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    bool valid = i < n;
    int un = valid ? u[i] : -1, vn = valid ? v[i] : -1;
    valid = valid && (uint32_t)un < nnodes && (uint32_t)vn < nnodes;       // ids outside [0, nnodes) touch nothing
    unsigned long long b = valid ? (unsigned long long)__double_as_longlong(w[i]) : 0ull;
    if (valid) atomicMax(&nmax[vn], b);          // fast_scaled_scores.py:47-57
    unsigned act = __ballot_sync(0xffffffffu, valid);
    while (act) {
        const int l = __ffs(act) - 1;
        const int node = __shfl_sync(0xffffffffu, un, l);
        const bool mine = valid && un == node;
        const unsigned long long mx = warp_max_u64(mine ? b : 0ull);
        if (lane == l) atomicMax(&nmax[node], mx);
        act &= ~__ballot_sync(0xffffffffu, mine);
    }
}

__global__ void node_second_kernel(const int32_t *u, const int32_t *v, const double *w, size_t n, uint32_t nnodes,
                                   const unsigned long long *nmax, unsigned long long *nsecond, unsigned int *ntop)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    bool valid = i < n;
    int un = valid ? u[i] : -1, vn = valid ? v[i] : -1;
    valid = valid && (uint32_t)un < nnodes && (uint32_t)vn < nnodes;
    unsigned long long b = valid ? (unsigned long long)__double_as_longlong(w[i]) : 0ull;
    if (valid) {
        if (b == nmax[vn]) atomicAdd(&ntop[vn], 1u);
        else atomicMax(&nsecond[vn], b);
    }
    unsigned act = __ballot_sync(0xffffffffu, valid);
    while (act) {
        const int l = __ffs(act) - 1;
        const int node = __shfl_sync(0xffffffffu, un, l);
        const bool mine = valid && un == node;
        const unsigned long long top = nmax[node];
        const unsigned cnt = __popc(__ballot_sync(0xffffffffu, mine && b == top));
        const unsigned long long sec = warp_max_u64((mine && b != top) ? b : 0ull);
        if (lane == l) {
            if (cnt) atomicAdd(&ntop[node], cnt);
            if (sec) atomicMax(&nsecond[node], sec);
        }
        act &= ~__ballot_sync(0xffffffffu, mine);
    }
}

__global__ void ratio_kernel(const int32_t *u, const int32_t *v, const double *w, size_t n, uint32_t nnodes, const unsigned long long *nmax,
                             const unsigned long long *nsecond, const unsigned int *ntop, double *best_alt,
                             uint8_t *is_int1, double *ratio, int64_t *digits, int32_t *exp10, int64_t *bdigits,
                             int32_t *bexp10)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double wi = w[i];
    int a = u[i], b = v[i];
    if ((uint32_t)a >= nnodes || (uint32_t)b >= nnodes) {                  // documented precondition violated: a defined, inert result
        best_alt[i] = 0.0; is_int1[i] = 0; ratio[i] = 0.0; digits[i] = 0; exp10[i] = 0; bdigits[i] = 0; bexp10[i] = 0;
        return;
    }
    double mu = __longlong_as_double((long long)nmax[a]), mv = __longlong_as_double((long long)nmax[b]);
    double best = mu > mv ? mu : mv;                       // :79
    if (best == wi) {                                      // :80-82  this edge is the best at both ends
        // max over the OTHER edges at either end (get_max_weight, :13-33): the
        // second-largest weight with multiplicity
        double au = __longlong_as_double((long long)(ntop[a] >= 2 ? nmax[a] : nsecond[a]));
        double av = __longlong_as_double((long long)(ntop[b] >= 2 ? nmax[b] : nsecond[b]));
        best = au > av ? au : av;
    }
    uint8_t int1 = 0;
    if (best == 0.0) { best = 1.0; int1 = 1; }             // :83-84  the int literal 1
    best_alt[i] = best;
    is_int1[i] = int1;
    double r = __ddiv_rn(wi, best);                        // :86
    ratio[i] = r;
    double dummy;
    g12_any(r, &digits[i], &exp10[i], &dummy);
    g12_any(best, &bdigits[i], &bexp10[i], &dummy);     // text of column 4
}

__global__ void zero_nodes_kernel(unsigned long long *p, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = 0;
}

inline unsigned nblk(size_t n, int t) { return (unsigned)((n + t - 1) / t); }

}   // namespace

extern "C" int salsa_g12(const double *d_x, size_t n, int64_t *d_digits, int32_t *d_exp10, double *d_w12, void *stream)
{
    if (n == 0) return SALSA_OK;
    if (!d_x || !d_digits || !d_exp10 || !d_w12) SALSA_FAIL(SALSA_ERR_ARG, "salsa_g12: NULL argument");
    g12_kernel<<<nblk(n, 128), 128, 0, as_stream(stream)>>>(d_x, n, d_digits, d_exp10, d_w12);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}

extern "C" int salsa_links_score(const uint64_t *d_keys, const uint64_t *d_counts, size_t n, int32_t ncontig,
                                 const int64_t *d_re_left, const int64_t *d_re_right, double *d_score, uint8_t *d_is_int0,
                                 int64_t *d_digits, int32_t *d_exp10, double *d_w12, int32_t *d_u, int32_t *d_v, void *stream)
{
    if (n == 0) return SALSA_OK;
    if (!d_keys || !d_counts || !d_re_left || !d_re_right || !d_score || !d_is_int0 || !d_digits || !d_exp10 || !d_w12 ||
        !d_u || !d_v || ncontig <= 0)
        SALSA_FAIL(SALSA_ERR_ARG, "salsa_links_score: bad argument");
    score_kernel<<<nblk(n, 128), 128, 0, as_stream(stream)>>>(d_keys, d_counts, n, (uint32_t)ncontig, d_re_left, d_re_right,
                                                              d_score, d_is_int0, d_digits, d_exp10, d_w12, d_u, d_v);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}

extern "C" size_t salsa_scaled_scratch_bytes(int32_t nnodes) { return (size_t)(nnodes > 0 ? nnodes : 0) * 24 + 64; }

// scratch layout: nmax u64[nnodes] | nsecond u64[nnodes] | ntop u32[nnodes] (+pad)
static int scaled_launch(const int32_t *d_u, const int32_t *d_v, const double *d_w, size_t n, int32_t nnodes,
                         unsigned long long *nmax, double *d_best_alt, uint8_t *d_is_int1, double *d_ratio,
                         int64_t *d_digits, int32_t *d_exp10, int64_t *d_bdigits, int32_t *d_bexp10, cudaStream_t st)
{
    unsigned long long *nsecond = nmax + nnodes;
    unsigned int *ntop = (unsigned int *)(nsecond + nnodes);
    size_t words = (size_t)nnodes * 2 + ((size_t)nnodes + 1) / 2;
    zero_nodes_kernel<<<nblk(words, 256), 256, 0, st>>>(nmax, words);
    node_max_kernel<<<nblk(n, 256), 256, 0, st>>>(d_u, d_v, d_w, n, (uint32_t)nnodes, nmax);
    node_second_kernel<<<nblk(n, 256), 256, 0, st>>>(d_u, d_v, d_w, n, (uint32_t)nnodes, nmax, nsecond, ntop);
    ratio_kernel<<<nblk(n, 128), 128, 0, st>>>(d_u, d_v, d_w, n, (uint32_t)nnodes, nmax, nsecond, ntop, d_best_alt, d_is_int1, d_ratio,
                                               d_digits, d_exp10, d_bdigits, d_bexp10);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}

extern "C" int salsa_scaled_scores(const int32_t *d_u, const int32_t *d_v, const double *d_w, size_t n, int32_t nnodes,
                                   void *d_scratch, double *d_best_alt, uint8_t *d_is_int1, double *d_ratio,
                                   int64_t *d_ratio_digits, int32_t *d_ratio_exp10, int64_t *d_best_digits,
                                   int32_t *d_best_exp10, void *stream)
{
    if (n == 0) return SALSA_OK;
    if (!d_u || !d_v || !d_w || !d_scratch || !d_best_alt || !d_is_int1 || !d_ratio || !d_ratio_digits || !d_ratio_exp10 ||
        !d_best_digits || !d_best_exp10 || nnodes <= 0)
        SALSA_FAIL(SALSA_ERR_ARG, "salsa_scaled_scores: bad argument");
    return scaled_launch(d_u, d_v, d_w, n, nnodes, (unsigned long long *)d_scratch, d_best_alt, d_is_int1, d_ratio,
                         d_ratio_digits, d_ratio_exp10, d_best_digits, d_best_exp10, as_stream(stream));
}

// ---- best-buddy ratios for callers that must not allocate -----------------
// Some hosts of this library cannot obtain device memory at call time: a
// plugin running inside another program's stream-ordered section, a caller
// whose memory pool is frozen for the duration of a pipeline stage, or a
// thin FFI binding (cgo / JNI style) that is only handed input and output
// buffers and has no allocator of its own to produce a scratch buffer.  For
// them the per-node working arrays live in this fixed device array instead of
// a caller-provided buffer.  It is part of the loaded module, so every GPU
// that loads libsalsa_b200 holds these bytes until the process exits.
#define SALSA_STATIC_MAX_NODES 65536
static __device__ __align__(256) unsigned char g_scaled_scratch[(size_t)SALSA_STATIC_MAX_NODES * 24 + 256];

// In host code `g_scaled_scratch` is only a handle for the symbol, not a
// device address; the address a kernel can use is looked up once per device.
static std::mutex g_static_mu;
static void *g_static_ptr[64];

static int static_scratch_ptr(void **out)
{
    int dev = -1;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) SALSA_FAIL(SALSA_ERR_UNSUPPORTED, "device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[dev]) {
        void *p = NULL;
        CUDA_TRY(cudaGetSymbolAddress(&p, g_scaled_scratch));   // resolved on the CURRENT device
        if (!p) SALSA_FAIL(SALSA_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
        g_static_ptr[dev] = p;
    }
    *out = g_static_ptr[dev];
    return SALSA_OK;
}

extern "C" int32_t salsa_scaled_scores_static_max_nodes(void) { return SALSA_STATIC_MAX_NODES; }

extern "C" int salsa_scaled_scores_static(const int32_t *d_u, const int32_t *d_v, const double *d_w, size_t n,
                                          int32_t nnodes, double *d_best_alt, uint8_t *d_is_int1, double *d_ratio,
                                          int64_t *d_ratio_digits, int32_t *d_ratio_exp10, int64_t *d_best_digits,
                                          int32_t *d_best_exp10, void *stream)
{
    if (nnodes > SALSA_STATIC_MAX_NODES)
        SALSA_FAIL(SALSA_ERR_ARG,
                   "salsa_scaled_scores_static: %d nodes exceed the built-in scratch limit of %d nodes "
                   "(use salsa_scaled_scores with a caller-provided scratch buffer)",
                   nnodes, SALSA_STATIC_MAX_NODES);
    if (n == 0) return SALSA_OK;
    if (!d_u || !d_v || !d_w || !d_best_alt || !d_is_int1 || !d_ratio || !d_ratio_digits || !d_ratio_exp10 ||
        !d_best_digits || !d_best_exp10 || nnodes <= 0)
        SALSA_FAIL(SALSA_ERR_ARG, "salsa_scaled_scores_static: bad argument");
    void *scratch = NULL;
    int rc = static_scratch_ptr(&scratch);
    if (rc) return rc;
    // scaled_launch zeroes the node arrays it uses before the first atomic, so
    // whatever the previous call left behind is never read.
    return scaled_launch(d_u, d_v, d_w, n, nnodes, (unsigned long long *)scratch, d_best_alt, d_is_int1, d_ratio,
                         d_ratio_digits, d_ratio_exp10, d_best_digits, d_best_exp10, as_stream(stream));
}

// ---- host text assembly from (digits, exp10): Python-2 str(float) ---------
static int fmt_g12(int64_t digits, int32_t k, char *out)
{
    char *p = out;
    if (digits == 0) { memcpy(p, "0.0", 3); return 3; }
    if (digits == SALSA_G12_INF) { memcpy(p, "inf", 3); return 3; }          // Python-2 str(float('inf'))
    if (digits == SALSA_G12_NINF) { memcpy(p, "-inf", 4); return 4; }
    if (digits == SALSA_G12_NAN) { memcpy(p, "nan", 3); return 3; }
    if (digits == SALSA_G12_RANGE) return -1;                                 // |x| outside [1e-30, 1e30]: no digits were computed
    if (digits < 0) { *p++ = '-'; digits = -digits; }
    char d[16];
    int nd = 12;
    for (int i = 11; i >= 0; i--) { d[i] = (char)('0' + digits % 10); digits /= 10; }
    while (nd > 1 && d[nd - 1] == '0') nd--;          // %g strips trailing zeros
    if (k < -4 || k >= 12) {                          // scientific
        *p++ = d[0];
        if (nd > 1) { *p++ = '.'; memcpy(p, d + 1, nd - 1); p += nd - 1; }
        *p++ = 'e';
        *p++ = k < 0 ? '-' : '+';
        int a = k < 0 ? -k : k;
        if (a >= 100) { *p++ = (char)('0' + a / 100); a %= 100; }
        *p++ = (char)('0' + a / 10);
        *p++ = (char)('0' + a % 10);
    } else if (k >= 0) {
        int ip = k + 1;                               // digits before the point
        for (int i = 0; i < ip; i++) *p++ = i < nd ? d[i] : '0';
        *p++ = '.';
        if (nd > ip) { memcpy(p, d + ip, nd - ip); p += nd - ip; }
        else *p++ = '0';                              // Python's '.0' for integral text
    } else {
        *p++ = '0'; *p++ = '.';
        for (int i = 0; i < -k - 1; i++) *p++ = '0';
        memcpy(p, d, nd); p += nd;
    }
    return (int)(p - out);
}

static inline char *put_i64(char *p, int64_t v)
{
    char t[24];
    int n = 0;
    bool neg = v < 0;
    uint64_t a = neg ? (uint64_t)(-v) : (uint64_t)v;
    do { t[n++] = (char)('0' + a % 10); a /= 10; } while (a);
    if (neg) *p++ = '-';
    while (n) *p++ = t[--n];
    return p;
}

extern "C" int salsa_format_g12(const int64_t *h_digits, const int32_t *h_exp10, size_t n, char *h_out, size_t out_cap,
                                int64_t *h_off)
{
    size_t pos = 0;
    for (size_t i = 0; i < n; i++) {
        if (pos + 32 > out_cap) SALSA_FAIL(SALSA_ERR_ARG, "salsa_format_g12: output buffer too small");
        h_off[i] = (int64_t)pos;
        int len = fmt_g12(h_digits[i], h_exp10[i], h_out + pos);
        if (len < 0) SALSA_FAIL(SALSA_ERR_UNSUPPORTED, "salsa_format_g12: value %zu lies outside [1e-30, 1e30]", i);
        pos += (size_t)len;
    }
    h_off[n] = (int64_t)pos;
    return SALSA_OK;
}

// The two file formatters below run on all host threads: lines are independent, so the edge range is cut into one
// contiguous piece per thread, every piece is written into its own buffer, and the pieces are copied behind one another
// (11 M link lines took 0.4 s per call on one thread -- the largest single item of a pipeline iteration's first pass).
// fmt(lo, hi, dst, cap) -> bytes written, or -(index + 1) of the first edge it cannot write / -(n + 1) for "buffer too small".
template <class F>
static int64_t format_parallel(size_t n, char *out, size_t out_cap, size_t line_bound, const F &fmt, const char *who)
{
    unsigned hw = std::thread::hardware_concurrency();
    size_t T = hw ? hw : 1;
    if (T > 32) T = 32;
    if (T > n / 32768 + 1) T = n / 32768 + 1;
    if (T <= 1) {
        int64_t r = fmt(0, n, out, out_cap);
        if (r == -(int64_t)(n + 1)) { salsa_set_error("%s: buffer too small", who); return -1; }
        if (r < 0) { salsa_set_error("%s: a value of edge %zu lies outside [1e-30, 1e30]", who, (size_t)(-r - 1)); return -1; }
        return r;
    }
    std::vector<std::unique_ptr<char[]>> buf(T);        // not value-initialised: the pieces are written, never read before
    std::vector<int64_t> got(T, 0);
    std::vector<std::thread> th;
    for (size_t t = 0; t < T; t++)
        th.emplace_back([&, t]() {
            const size_t lo = n * t / T, hi = n * (t + 1) / T;
            const size_t cap = (hi - lo) * line_bound + 64;
            buf[t].reset(new char[cap]);
            got[t] = fmt(lo, hi, buf[t].get(), cap);
        });
    for (auto &x : th) x.join();
    size_t total = 0;
    for (size_t t = 0; t < T; t++) {
        if (got[t] == -(int64_t)(n + 1)) { salsa_set_error("%s: buffer too small", who); return -1; }
        if (got[t] < 0) { salsa_set_error("%s: a value of edge %zu lies outside [1e-30, 1e30]", who, (size_t)(-got[t] - 1)); return -1; }
        total += (size_t)got[t];
    }
    if (total > out_cap) { salsa_set_error("%s: buffer too small", who); return -1; }
    th.clear();
    size_t pos = 0;
    for (size_t t = 0; t < T; t++) {
        th.emplace_back([&, t, pos]() { memcpy(out + pos, buf[t].get(), (size_t)got[t]); });
        pos += (size_t)got[t];
    }
    for (auto &x : th) x.join();
    return (int64_t)total;
}

// contig_links_iteration_N: "first:X\tsecond:Y\tscore\tcount\n" (make_links.py:169)
extern "C" int64_t salsa_format_links(const char *names_blob, const int64_t *name_off, const int64_t *first,
                                      const int64_t *second, const uint8_t *ends, const int64_t *count,
                                      const int64_t *digits, const int32_t *exp10, const uint8_t *is_int0, size_t n,
                                      char *out, size_t out_cap)
{
    size_t longest = 0;
    for (size_t i = 0; i < n; i++) {
        const size_t l = (size_t)(name_off[first[i] + 1] - name_off[first[i]]) + (size_t)(name_off[second[i] + 1] - name_off[second[i]]);
        if (l > longest) longest = l;
    }
    auto fmt = [&](size_t lo, size_t hi, char *dst, size_t cap) -> int64_t {
        char *p = dst;
        for (size_t i = lo; i < hi; i++) {
            int64_t a = first[i], b = second[i];
            size_t la = (size_t)(name_off[a + 1] - name_off[a]), lb = (size_t)(name_off[b + 1] - name_off[b]);
            if ((size_t)(p - dst) + la + lb + 80 > cap) return -(int64_t)(n + 1);
            memcpy(p, names_blob + name_off[a], la); p += la;
            *p++ = ':'; *p++ = (ends[i] & 2) ? 'E' : 'B'; *p++ = '\t';
            memcpy(p, names_blob + name_off[b], lb); p += lb;
            *p++ = ':'; *p++ = (ends[i] & 1) ? 'E' : 'B'; *p++ = '\t';
            if (is_int0[i]) *p++ = '0';
            else {
                int len = fmt_g12(digits[i], exp10[i], p);
                if (len < 0) return -(int64_t)(i + 1);
                p += len;
            }
            *p++ = '\t';
            p = put_i64(p, count[i]);
            *p++ = '\n';
        }
        return (int64_t)(p - dst);
    };
    return format_parallel(n, out, out_cap, longest + 80, fmt, "salsa_format_links");
}

// contig_links_scaled_iteration_N (fast_scaled_scores.py:86):
// u v str(w) str(bestAlt) str(w/bestAlt) links False False
extern "C" int64_t salsa_format_scaled(const char *node_blob, const int64_t *node_off, const int32_t *u, const int32_t *v,
                                       const int64_t *w_digits, const int32_t *w_exp10, const int64_t *b_digits,
                                       const int32_t *b_exp10, const uint8_t *b_is_int1, const int64_t *r_digits,
                                       const int32_t *r_exp10, const int64_t *links, size_t n, char *out, size_t out_cap)
{
    size_t longest = 0;
    for (size_t i = 0; i < n; i++) {
        const size_t l = (size_t)(node_off[u[i] + 1] - node_off[u[i]]) + (size_t)(node_off[v[i] + 1] - node_off[v[i]]);
        if (l > longest) longest = l;
    }
    auto fmt = [&](size_t lo, size_t hi, char *dst, size_t cap) -> int64_t {
        char *p = dst;
        for (size_t i = lo; i < hi; i++) {
            int64_t a = u[i], b = v[i];
            size_t la = (size_t)(node_off[a + 1] - node_off[a]), lb = (size_t)(node_off[b + 1] - node_off[b]);
            if ((size_t)(p - dst) + la + lb + 140 > cap) return -(int64_t)(n + 1);
            memcpy(p, node_blob + node_off[a], la); p += la; *p++ = '\t';
            memcpy(p, node_blob + node_off[b], lb); p += lb; *p++ = '\t';
            int l1 = fmt_g12(w_digits[i], w_exp10[i], p);
            if (l1 >= 0) { p += l1; *p++ = '\t'; }
            int l2 = b_is_int1[i] ? 1 : fmt_g12(b_digits[i], b_exp10[i], p);
            if (b_is_int1[i]) *p = '1';
            if (l2 >= 0) { p += l2; *p++ = '\t'; }
            int l3 = fmt_g12(r_digits[i], r_exp10[i], p);
            if (l1 < 0 || l2 < 0 || l3 < 0) return -(int64_t)(i + 1);
            p += l3; *p++ = '\t';
            p = put_i64(p, links[i]);
            memcpy(p, "\tFalse\tFalse\n", 13); p += 13;
        }
        return (int64_t)(p - dst);
    };
    return format_parallel(n, out, out_cap, longest + 140, fmt, "salsa_format_scaled");
}
