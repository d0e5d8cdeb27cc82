// links_table.cu -- link table -> ordered edge list.
//
// Replaces the dict walk that writes contig_links_iteration_N
// (make_links.py:161-169).  The reference's line order is Python-2 dict order,
// which no consumer depends on; this library emits edges in ascending key
// order = (first name, second name, ends) order.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace {

constexpr int CT = 256;            // threads per block
constexpr int CW = CT / 32;
constexpr int CHUNK = 8192;        // cells per block
constexpr int WCHUNK = CHUNK / CW; // cells per warp (contiguous)

__device__ __forceinline__ bool live(const uint64_t *keys, const uint64_t *counts, size_t i)
{
    return keys ? (keys[i] != ~0ull && counts[i] != 0) : (counts[i] != 0);
}

__global__ void __launch_bounds__(CT) count_kernel(const uint64_t *keys, const uint64_t *counts, size_t n, uint64_t *blk)
{
    size_t base = (size_t)blockIdx.x * CHUNK;
    int c = 0;
    for (int i = threadIdx.x; i < CHUNK; i += CT) {
        size_t j = base + i;
        if (j < n && live(keys, counts, j)) c++;
    }
    __shared__ int red[CW];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < CW; w++) t += red[w];
        blk[blockIdx.x] = (uint64_t)t;
    }
}

// single block: exclusive scan of blk[0..nb) in place, total -> blk[nb]
__global__ void __launch_bounds__(1024) scan_kernel(uint64_t *blk, size_t nb)
{
    __shared__ uint64_t wsum[32];
    __shared__ uint64_t carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (size_t base = 0; base < nb; base += 1024) {
        size_t i = base + threadIdx.x;
        uint64_t v = i < nb ? blk[i] : 0, x = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint64_t o = __shfl_up_sync(0xffffffffu, x, d);
            if ((threadIdx.x & 31) >= d) x += o;
        }
        if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = x;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint64_t w = wsum[threadIdx.x], y = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                uint64_t o = __shfl_up_sync(0xffffffffu, y, d);
                if (threadIdx.x >= d) y += o;
            }
            wsum[threadIdx.x] = y - w;
        }
        __syncthreads();
        uint64_t excl = carry_s + wsum[threadIdx.x >> 5] + x - v;
        if (i < nb) blk[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) blk[nb] = carry_s;
}

__global__ void __launch_bounds__(CT) write_kernel(const uint64_t *keys, const uint64_t *counts, size_t n,
                                                   const uint64_t *blk, uint64_t *okeys, uint64_t *ocounts, size_t ocap)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    size_t wbase = (size_t)blockIdx.x * CHUNK + (size_t)warp * WCHUNK;
    int c = 0;
    for (int i = 0; i < WCHUNK; i += 32) {
        size_t j = wbase + i + lane;
        bool l = j < n && live(keys, counts, j);
        c += __popc(__ballot_sync(0xffffffffu, l));
    }
    __shared__ int wc[CW];
    if (lane == 0) wc[warp] = c;
    __syncthreads();
    uint64_t off = blk[blockIdx.x];
    for (int w = 0; w < warp; w++) off += (uint64_t)wc[w];
    for (int i = 0; i < WCHUNK; i += 32) {
        size_t j = wbase + i + lane;
        bool l = j < n && live(keys, counts, j);
        unsigned m = __ballot_sync(0xffffffffu, l);
        if (l) {
            uint64_t o = off + (uint64_t)__popc(m & ((1u << lane) - 1u));
            if (o < ocap) {
                okeys[o] = keys ? keys[j] : (uint64_t)j;
                ocounts[o] = counts[j];
            }
        }
        off += (uint64_t)__popc(m);
    }
}

}   // namespace

// in-place exclusive scan of d_vals[0..n); the total lands in d_vals[n]
extern "C" int salsa_exclusive_scan_u64(uint64_t *d_vals, size_t n, void *stream)
{
    if (!d_vals) SALSA_FAIL(SALSA_ERR_ARG, "salsa_exclusive_scan_u64: NULL argument");
    scan_kernel<<<1, 1024, 0, as_stream(stream)>>>(d_vals, n);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}

extern "C" size_t salsa_links_compact_scratch_bytes(size_t ncells)
{
    size_t nb = (ncells + CHUNK - 1) / CHUNK;
    return (nb + 2) * sizeof(uint64_t);
}

extern "C" int salsa_links_compact(const uint64_t *d_keys, const uint64_t *d_counts, size_t ncells, uint64_t *d_out_keys,
                                   uint64_t *d_out_counts, size_t out_capacity, void *d_scratch, uint64_t *h_n,
                                   void *stream)
{
    if (!d_counts || !d_scratch || !h_n) SALSA_FAIL(SALSA_ERR_ARG, "salsa_links_compact: NULL argument");
    cudaStream_t st = as_stream(stream);
    size_t nb = (ncells + CHUNK - 1) / CHUNK;
    uint64_t *blk = (uint64_t *)d_scratch;
    *h_n = 0;
    if (nb == 0) return SALSA_OK;
    if (nb > 0x7fffffffull) SALSA_FAIL(SALSA_ERR_ARG, "table too large");
    count_kernel<<<(unsigned)nb, CT, 0, st>>>(d_keys, d_counts, ncells, blk);
    scan_kernel<<<1, 1024, 0, st>>>(blk, nb);
    CUDA_TRY(cudaGetLastError());
    uint64_t total = 0;
    CUDA_TRY(cudaMemcpyAsync(&total, blk + nb, sizeof total, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    *h_n = total;
    if (!d_out_keys || !d_out_counts) return SALSA_OK;      // count only
    if (total > out_capacity) SALSA_FAIL(SALSA_ERR_ARG, "salsa_links_compact: %llu links, capacity %zu",
                                         (unsigned long long)total, out_capacity);
    write_kernel<<<(unsigned)nb, CT, 0, st>>>(d_keys, d_counts, ncells, blk, d_out_keys, d_out_counts, out_capacity);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}

// key-sorts (key, count) pairs; used after compacting a sparse table.  The
// radix sort is CUB's (library code: this step is off the hot path).
extern "C" int salsa_sort_pairs_u64(uint64_t *d_keys_in, uint64_t *d_vals_in, uint64_t *d_keys_out, uint64_t *d_vals_out,
                                    size_t n, void *d_tmp, size_t *tmp_bytes, void *stream)
{
    if (!tmp_bytes) SALSA_FAIL(SALSA_ERR_ARG, "tmp_bytes is NULL");
    if (n > 0x7fffffffull) SALSA_FAIL(SALSA_ERR_ARG, "too many pairs to sort");
    size_t need = 0;
    CUDA_TRY(cub::DeviceRadixSort::SortPairs(NULL, need, d_keys_in, d_keys_out, d_vals_in, d_vals_out, (int)n, 0, 64,
                                             as_stream(stream)));
    if (!d_tmp) { *tmp_bytes = need; return SALSA_OK; }
    if (*tmp_bytes < need) SALSA_FAIL(SALSA_ERR_ARG, "sort scratch too small: need %zu", need);
    CUDA_TRY(cub::DeviceRadixSort::SortPairs(d_tmp, need, d_keys_in, d_keys_out, d_vals_in, d_vals_out, (int)n, 0, 64,
                                             as_stream(stream)));
    return SALSA_OK;
}
