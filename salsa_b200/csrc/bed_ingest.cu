// bed_ingest.cu -- BED text -> contig-end link histogram, one pass over HBM.
//
// Replaces the per-line loop of the reference's make_links.py:60-146
// (str.split, adjacent-line pairing, read.split('/')[0] test, B/E end
// classification, contig_links[key] += 1).  The tile / line / record machinery
// is bed_scan.cuh; this file supplies what an adjacent pair contributes:
// one 64-bit RED into the dense table, or one insert into the sparse table.
// HBM traffic = the BED bytes once; everything else lives in smem / L2.
#include <stdlib.h>

#include <algorithm>
#include <string>
#include <vector>

#include "bed_scan.cuh"
#include "bed_stream.cuh"

namespace {

using namespace bedscan;

struct LinkArgs {
    uint64_t *table;          // dense: counts[n*n*4]; sparse: counts[capacity]
    uint64_t *keys;           // sparse only
    uint64_t cap_mask;        // sparse only
};

template <bool SPARSE, bool AVOID_>
struct LinkPolicy {
    typedef LinkArgs Args;
    static constexpr bool AVOID = AVOID_;   // links_avoid_iteration_N present (make_links.py:29-35,72-74)
    static constexpr bool NEED_RFULL = false;
#ifdef SALSA_NO_LAZY
    static constexpr bool LAZY = false;
#else
    // two-phase warp loop (bed_scan.cuh): only pairs whose contig names differ and whose read names may be
    // equal get the full parse; pair() below then sees exactly the records it always saw
    static constexpr bool LAZY = !AVOID_;
#endif
    static __device__ __forceinline__ bool lazy_candidate(bool same_name, bool reads_may_equal) { return !same_name && reads_may_equal; }
    // make_links.py:78 compares read.split('/')[0]
    static __device__ __forceinline__ int read_cmp_len(int rslash, int /*rfull*/) { return rslash; }

    static __device__ __forceinline__ void emit(const LinkArgs &L, const ScanArgs &A, uint64_t key, uint64_t line_off)
    {
        if (!SPARSE) {
            atomicAdd((unsigned long long *)&L.table[key], 1ull);
        } else {
            uint64_t slot = salsa_mix64(key) & L.cap_mask;
            for (uint64_t probe = 0; probe <= L.cap_mask; probe++) {
                unsigned long long k = ((volatile unsigned long long *)L.keys)[slot];
                if (k == ~0ull) {
                    unsigned long long old = atomicCAS((unsigned long long *)&L.keys[slot], ~0ull, (unsigned long long)key);
                    k = (old == ~0ull || old == key) ? key : old;
                }
                if (k == key) {
                    atomicAdd((unsigned long long *)&L.table[slot], 1ull);
                    return;
                }
                slot = (slot + 1) & L.cap_mask;
            }
            report(A.stats, line_off, SALSA_DATA_TABLE_FULL);
        }
    }

    // the reference's body for one adjacent-line pair: make_links.py:70-144
    static __device__ __forceinline__ bool pair(const LinkArgs &L, const ScanArgs &A, const uint8_t *s, const Rec &pv,
                                                const Rec &cu, uint64_t line_off, bool /*prev_is_first*/)
    {
        uint32_t pnf = pv.flags & 7u, cnf = cu.flags & 7u;
        if ((pv.flags | cu.flags) & 16u) { report(A.stats, line_off, SALSA_DATA_LINE_TOO_LONG); return false; }
        if (pnf < 1 || cnf < 1) { report(A.stats, line_off, SALSA_DATA_INDEX_ERROR); return false; }   // :71
        if (same_contig(s, pv, cu)) return false;
        if (pnf < 4 || cnf < 4) { report(A.stats, line_off, SALSA_DATA_INDEX_ERROR); return false; }   // :76-77
        if (!reads_equal(s, pv, cu)) return false;                                                     // :78
        if ((pv.flags | cu.flags) & 8u) { report(A.stats, line_off, SALSA_DATA_VALUE_ERROR); return false; }  // :81-82
        if (pv.id < 0 || cu.id < 0) { report(A.stats, line_off, SALSA_DATA_KEY_ERROR); return false; }  // :101-102
        ContigMeta mp = A.t.meta[pv.id], mc = A.t.meta[cu.id];
        if (!(mp.flags & mc.flags & SALSA_CTG_HAS_LEN)) { report(A.stats, line_off, SALSA_DATA_KEY_ERROR); return false; }
        if (mp.len_floor == 0 || mc.len_floor == 0) {
            // float(len) == 0 raises ZeroDivisionError (:106-107); lengths in (0,1) are refused on the host
            report(A.stats, line_off, SALSA_DATA_ZERO_DIV);
            return false;
        }
        if (!(mp.flags & mc.flags & SALSA_CTG_HAS_RE)) { report(A.stats, line_off, SALSA_DATA_KEY_ERROR); return false; }  // :113
        long long sp = rec_start(pv) + rec_end(pv), sc = rec_start(cu) + rec_end(cu);
        // B <=> pos <= len/2 <=> start+end <= floor(len); E otherwise (:108-139)
        uint32_t ep = sp > mp.len_floor, ec = sc > mc.len_floor;
        uint32_t first, second, ends;
        if (pv.id < cu.id) { first = pv.id; second = cu.id; ends = ep * 2 + ec; }     // :90-96
        else { first = cu.id; second = pv.id; ends = ec * 2 + ep; }
        uint64_t key = (((uint64_t)first * (uint64_t)A.t.n + second) << 2) | ends;
        emit(L, A, key, line_off);
        return true;
    }
};

// what a counted pair does to the table, for the streaming scanner (bed_stream.cuh)
template <bool SPARSE>
struct LinkEmit {
    typedef LinkArgs Args;
    static __device__ __forceinline__ void emit(const LinkArgs &L, const ScanArgs &A, uint64_t key, uint64_t line_off)
    {
        LinkPolicy<SPARSE, false>::emit(L, A, key, line_off);
    }
};

// SALSA_TILE_SCANNER=1 keeps the tile kernel of bed_scan.cuh for the plain link scan (A/B measurements)
static bool use_tile_scanner()
{
    static int v = -1;
    if (v < 0) {
        const char *e = getenv("SALSA_TILE_SCANNER");
        v = (e && *e && *e != '0') ? 1 : 0;
    }
    return v == 1;
}

template <bool SPARSE>
static int scan_links(const LinkArgs &L, salsa_ctx *ctx, const void *d_bed, size_t nbytes, uint64_t base_offset, uint64_t *d_stats,
                      cudaStream_t st)
{
    void *slot = NULL;
    int rc = salsa_timer_begin(ctx, st, &slot);            // salsa_timers: no-ops unless enabled
    if (rc) return rc;
    if (ctx->t.navoid > 0)
        rc = bedscan::launch_scan<LinkPolicy<SPARSE, true>>(L, ctx->t, d_bed, nbytes, base_offset, d_stats, st);
    else if (use_tile_scanner())
        rc = bedscan::launch_scan<LinkPolicy<SPARSE, false>>(L, ctx->t, d_bed, nbytes, base_offset, d_stats, st);
    else
        rc = bedstream::launch_stream<LinkEmit<SPARSE>>(L, ctx->t, d_bed, nbytes, base_offset, d_stats, ctx->scan_mode, ctx->names_punct,
                                                        ctx->d_mode, st);
    salsa_timer_end(ctx, st, slot);
    return rc;
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
    CUDA_TRY(cudaSetDevice(ctx->device));
    return SALSA_OK;
}

extern "C" int salsa_bed_ingest_dense(salsa_ctx *ctx, const void *d_bed, size_t nbytes, uint64_t base_offset,
                                      uint64_t *d_table, uint64_t *d_stats, void *stream)
{
    int rc = check_ingest_args(ctx, d_bed, nbytes, d_table, d_stats);
    if (rc) return rc;
    LinkArgs L{d_table, NULL, 0};
    return scan_links<false>(L, ctx, d_bed, nbytes, base_offset, d_stats, as_stream(stream));
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
    LinkArgs L{d_counts, d_keys, capacity - 1};
    return scan_links<true>(L, ctx, d_bed, nbytes, base_offset, d_stats, as_stream(stream));
}

// Is the line that starts at p (and ends before lim) on a contig that occurs in some avoid pair?  (str.split()'s first token)
static bool line_involved(const salsa_ctx *ctx, const char *p, const char *lim)
{
    const std::vector<std::string> *inv = (const std::vector<std::string> *)ctx->h_involved;
    if (!inv || inv->empty()) return false;
    auto ws = [](char c) { return c == ' ' || (c >= 9 && c <= 13); };
    while (p < lim && *p != '\n' && ws(*p)) p++;
    const char *e = p;
    while (e < lim && !ws(*e)) e++;
    return std::binary_search(inv->begin(), inv->end(), std::string(p, (size_t)(e - p)));
}

// Host-resident BED: chunked H2D copies overlapped with the kernel.  Chunks
// start on line starts and overlap by exactly one line (see salsa_b200.h); with
// an avoid set, on a line whose contig is in no avoid pair (it can never be skipped).
extern "C" int salsa_bed_ingest_dense_host(salsa_ctx *ctx, const void *h_bed, size_t nbytes, void *d_stage,
                                           size_t chunk_bytes, uint64_t *d_table, uint64_t *d_stats, void *stream)
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
    LinkArgs L{d_table, NULL, 0};
    // the caller's earlier work on d_table / d_stats (zeroing) was enqueued on `stream`: both internal streams follow it
    CUDA_TRY(cudaEventRecord(ctx->ev_user, as_stream(stream)));
    CUDA_TRY(cudaStreamWaitEvent(ctx->s_copy, ctx->ev_user, 0));
    CUDA_TRY(cudaStreamWaitEvent(ctx->s_comp, ctx->ev_user, 0));
    while (pos < nbytes) {
        size_t end = pos + chunk_bytes < nbytes ? pos + chunk_bytes : nbytes;
        size_t next = nbytes;
        if (end < nbytes) {
            // the last complete line inside [pos, end) is shared: it ends this chunk and starts the next one
            const char *nl = (const char *)memrchr(h + pos, '\n', end - pos);
            const char *pl = nl && nl > h + pos ? (const char *)memrchr(h + pos, '\n', (size_t)(nl - (h + pos))) : NULL;
            while (pl && ctx->t.navoid > 0 && line_involved(ctx, pl + 1, nl)) {
                // the shared line could be skipped by the avoid rule: move the cut back to a line that cannot
                nl = pl;
                pl = nl > h + pos ? (const char *)memrchr(h + pos, '\n', (size_t)(nl - (h + pos))) : NULL;
            }
            if (!nl || !pl)
                SALSA_FAIL(SALSA_ERR_ARG, ctx->t.navoid > 0 ? "no line outside the avoid list to split the BED at near byte %zu (or a line longer than the staging chunk)"
                                                             : "BED line longer than the staging chunk near byte %zu", pos);
            end = (size_t)(nl - h) + 1;
            next = (size_t)(pl - h) + 1;      // the last line is the next chunk's "previous"
        }
        char *dst = (char *)d_stage + (size_t)(k & 1) * chunk_bytes;
        CUDA_TRY(cudaStreamWaitEvent(ctx->s_copy, ctx->ev_comp[k & 1], 0));   // buffer free again
        CUDA_TRY(cudaMemcpyAsync(dst, h + pos, end - pos, cudaMemcpyHostToDevice, ctx->s_copy));
        CUDA_TRY(cudaEventRecord(ctx->ev_copy[k & 1], ctx->s_copy));
        CUDA_TRY(cudaStreamWaitEvent(ctx->s_comp, ctx->ev_copy[k & 1], 0));
        rc = scan_links<false>(L, ctx, dst, end - pos, pos, d_stats, ctx->s_comp);
        if (rc) return rc;
        CUDA_TRY(cudaEventRecord(ctx->ev_comp[k & 1], ctx->s_comp));
        pos = next;
        k++;
    }
    // what the caller enqueues on `stream` next sees the finished table; the host buffer is free when this returns
    CUDA_TRY(cudaEventRecord(ctx->ev_user, ctx->s_comp));
    CUDA_TRY(cudaStreamWaitEvent(as_stream(stream), ctx->ev_user, 0));
    CUDA_TRY(cudaStreamSynchronize(ctx->s_comp));
    return SALSA_OK;
}
