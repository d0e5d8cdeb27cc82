// clean_bed.cu -- the BED rewrite of the cleaning step (-m yes).
//
// Replaces the streaming loop of correct.py:62-92: every line of the input BED is split with str.split(); a line on a
// contig that break_contigs_start did NOT cut is copied as its first four fields; a line on a cut contig is dropped when
// it straddles the cut (start < cut < end), keeps its coordinates under the name `contig_1` when it lies left of it
// (end < cut), and moves to `contig_2` with both coordinates shifted by the cut otherwise.  (The reference's other quirk
// -- the last, partial buffer of a million lines is never written -- is applied by the caller, which knows the ranks.)
//
// One thread per line, the line read from global memory (this runs once per pipeline, over the BED that is resident in
// HBM anyway): a PLAN pass returns the length of what each line will write (0 = dropped) and the first error, the caller
// turns the lengths into offsets, an EMIT pass writes.  HBM traffic: the BED twice, the output once.
#include "common.cuh"

namespace {

__device__ __forceinline__ bool ws(uint32_t c) { return c == 32u || (c - 9u) <= 4u; }

struct Tok {
    unsigned long long off;
    uint32_t len;
};

// name -> id by the context's table (exact: names up to 16 bytes from the slot, longer ones verified against the blob)
__device__ int lookup(const DevTable &t, const uint8_t *g, const Tok &nm)
{
    uint32_t w[4] = {0, 0, 0, 0};
    for (uint32_t i = 0; i < nm.len && i < 16; i++) w[i >> 2] |= (uint32_t)g[nm.off + i] << (8 * (i & 3));
    uint32_t h = salsa_name_hash(w[0], w[1], w[2], w[3], nm.len) & t.mask;
    for (;;) {
        const uint4 *sp = reinterpret_cast<const uint4 *>(&t.slots[h]);
        const uint4 k = __ldg(sp), m = __ldg(sp + 1);
        if (m.x == 0xFFFFFFFFu) return -1;
        if (m.x == nm.len && k.x == w[0] && k.y == w[1] && k.z == w[2] && k.w == w[3]) {
            bool eq = true;
            for (uint32_t i = 16; i < nm.len; i++)
                if ((uint8_t)__ldg(&t.blob[m.z + i]) != g[nm.off + i]) { eq = false; break; }
            if (eq) return (int)m.y;
        }
        h = (h + 1) & t.mask;
    }
}

// Python-2 int(token) for what a BED coordinate can be: [+-]?[0-9]{1,18}
__device__ bool to_int(const uint8_t *g, const Tok &t, long long &v)
{
    uint32_t i = 0;
    bool neg = false;
    if (t.len && (g[t.off] == '-' || g[t.off] == '+')) { neg = g[t.off] == '-'; i = 1; }
    if (t.len - i < 1 || t.len - i > 18) return false;
    unsigned long long a = 0;
    for (; i < t.len; i++) {
        uint32_t d = (uint32_t)g[t.off + i] - '0';
        if (d > 9u) return false;
        a = a * 10ull + d;
    }
    v = neg ? -(long long)a : (long long)a;
    return true;
}

__device__ __forceinline__ uint32_t ndigits(long long v)
{
    unsigned long long a = v < 0 ? (unsigned long long)(-v) : (unsigned long long)v;
    uint32_t n = 1;
    while (a >= 10ull) { a /= 10ull; n++; }
    return n + (v < 0 ? 1u : 0u);
}

__device__ __forceinline__ uint8_t *put_int(uint8_t *p, long long v)
{
    char t[24];
    int n = 0;
    unsigned long long a = v < 0 ? (unsigned long long)(-v) : (unsigned long long)v;
    do { t[n++] = (char)('0' + a % 10ull); a /= 10ull; } while (a);
    if (v < 0) *p++ = '-';
    while (n) *p++ = (uint8_t)t[--n];
    return p;
}

__device__ __forceinline__ uint8_t *put_tok(uint8_t *p, const uint8_t *g, const Tok &t)
{
    for (uint32_t i = 0; i < t.len; i++) p[i] = g[t.off + i];
    return p + t.len;
}

// What one line writes (out == NULL: only its length).  Errors: 1 = IndexError (< 4 fields, correct.py:67-69),
// 2 = ValueError (int() of a coordinate on a cut contig, :72-73), 3 = KeyError (contig not in the assembly, :69,81,86).
__device__ uint32_t clean_line(const DevTable &t, const int64_t *cut, const uint8_t *g, unsigned long long b, unsigned long long e,
                               uint8_t *out, int *err)
{
    Tok f[4];
    unsigned long long p = b;
    int nf = 0;
    while (nf < 4) {
        while (p < e && ws(g[p])) p++;
        if (p >= e) break;
        f[nf].off = p;
        while (p < e && !ws(g[p])) p++;
        f[nf].len = (uint32_t)(p - f[nf].off);
        nf++;
    }
    *err = 0;
    if (nf < 1) { *err = 1; return 0; }                  // attrs[0]
    const int id = lookup(t, g, f[0]);
    const long long bp = id >= 0 ? cut[id] : -1;
    if (id >= 0 && bp < 0) {                             // not cut: the first four fields as they are
        if (nf < 4) { *err = 1; return 0; }
        if (!out) return f[0].len + f[1].len + f[2].len + f[3].len + 4u;
        uint8_t *q = put_tok(out, g, f[0]); *q++ = '\t';
        q = put_tok(q, g, f[1]); *q++ = '\t';
        q = put_tok(q, g, f[2]); *q++ = '\t';
        q = put_tok(q, g, f[3]); *q++ = '\n';
        return (uint32_t)(q - out);
    }
    if (id < 0) {
        // `attrs[0] not in contig2breakpoints` holds for a name the assembly does not have either (the breakpoints file
        // names assembly contigs), and contig2new[attrs[0]] raises -- after attrs[1..3] were indexed? No: the lookup comes first
        *err = 3;
        return 0;
    }
    if (nf < 3) { *err = 1; return 0; }                  // int(attrs[1]), int(attrs[2])
    long long p1, p2;
    if (!to_int(g, f[1], p1) || !to_int(g, f[2], p2)) { *err = 2; return 0; }
    if (p1 < bp && p2 > bp) return 0;                    // straddles the cut: dropped
    if (nf < 4) { *err = 1; return 0; }                  // attrs[3]
    if (p2 < bp) {
        if (!out) return f[0].len + 2u + f[1].len + f[2].len + f[3].len + 4u;
        uint8_t *q = put_tok(out, g, f[0]); *q++ = '_'; *q++ = '1'; *q++ = '\t';
        q = put_tok(q, g, f[1]); *q++ = '\t';
        q = put_tok(q, g, f[2]); *q++ = '\t';
        q = put_tok(q, g, f[3]); *q++ = '\n';
        return (uint32_t)(q - out);
    }
    const long long s = p1 - bp, en = p2 - bp;
    if (!out) return f[0].len + 2u + ndigits(s) + ndigits(en) + f[3].len + 4u;
    uint8_t *q = put_tok(out, g, f[0]); *q++ = '_'; *q++ = '2'; *q++ = '\t';
    q = put_int(q, s); *q++ = '\t';
    q = put_int(q, en); *q++ = '\t';
    q = put_tok(q, g, f[3]); *q++ = '\n';
    return (uint32_t)(q - out);
}

__device__ __forceinline__ unsigned long long line_end(const uint8_t *g, unsigned long long b, unsigned long long nbytes)
{
    unsigned long long e = b;
    while (e < nbytes && g[e] != '\n') e++;
    return e;
}

__global__ void clean_plan_kernel(DevTable t, const int64_t *cut, const uint8_t *g, unsigned long long nbytes, const int64_t *starts,
                                  size_t nlines, uint32_t *out_len, unsigned long long *err)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nlines) return;
    const unsigned long long b = (unsigned long long)starts[i];
    int code;
    out_len[i] = clean_line(t, cut, g, b, line_end(g, b, nbytes), NULL, &code);
    if (code) atomicMin(err, ((unsigned long long)b << 8) | (unsigned long long)code);
}

__global__ void clean_emit_kernel(DevTable t, const int64_t *cut, const uint8_t *g, unsigned long long nbytes, const int64_t *starts,
                                  size_t nlines, const uint32_t *out_len, const int64_t *out_off, uint8_t *out)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nlines || out_len[i] == 0) return;
    const unsigned long long b = (unsigned long long)starts[i];
    int code;
    clean_line(t, cut, g, b, line_end(g, b, nbytes), out + out_off[i], &code);
}

}   // namespace

extern "C" int salsa_clean_bed_plan(salsa_ctx *ctx, const void *d_bed, size_t nbytes, const int64_t *d_cut, const int64_t *d_line_start,
                                    size_t nlines, uint32_t *d_out_len, uint64_t *d_err, void *stream)
{
    if (!ctx || (!d_bed && nbytes) || !d_cut || (!d_line_start && nlines) || (!d_out_len && nlines) || !d_err)
        SALSA_FAIL(SALSA_ERR_ARG, "salsa_clean_bed_plan: NULL argument");
    if (ctx->n <= 0) SALSA_FAIL(SALSA_ERR_ARG, "salsa_clean_bed_plan: call salsa_names_build first");
    CUDA_TRY(cudaSetDevice(ctx->device));
    cudaStream_t st = as_stream(stream);
    CUDA_TRY(cudaMemsetAsync(d_err, 0xFF, sizeof(uint64_t), st));
    if (nlines == 0) return SALSA_OK;
    clean_plan_kernel<<<(unsigned)((nlines + 127) / 128), 128, 0, st>>>(ctx->t, d_cut, (const uint8_t *)d_bed, nbytes, d_line_start, nlines,
                                                                       d_out_len, (unsigned long long *)d_err);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}

extern "C" int salsa_clean_bed_emit(salsa_ctx *ctx, const void *d_bed, size_t nbytes, const int64_t *d_cut, const int64_t *d_line_start,
                                    size_t nlines, const uint32_t *d_out_len, const int64_t *d_out_off, void *d_out, void *stream)
{
    if (!ctx || (!d_bed && nbytes) || !d_cut || (nlines && (!d_line_start || !d_out_len || !d_out_off || !d_out)))
        SALSA_FAIL(SALSA_ERR_ARG, "salsa_clean_bed_emit: NULL argument");
    if (ctx->n <= 0) SALSA_FAIL(SALSA_ERR_ARG, "salsa_clean_bed_emit: call salsa_names_build first");
    CUDA_TRY(cudaSetDevice(ctx->device));
    if (nlines == 0) return SALSA_OK;
    clean_emit_kernel<<<(unsigned)((nlines + 127) / 128), 128, 0, as_stream(stream)>>>(ctx->t, d_cut, (const uint8_t *)d_bed, nbytes,
                                                                                     d_line_start, nlines, d_out_len, d_out_off,
                                                                                     (uint8_t *)d_out);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}
