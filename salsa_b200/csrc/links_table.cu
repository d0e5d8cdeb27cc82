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

// ---- multi-GPU merge of sparse tables ---------------------------------------
// owner(key) = mix64(key) % nparts.  Links are bucketed by owner (histogram, scan,
// scatter), exchanged with one all-to-all, and added into the receiver's table.
namespace {

constexpr int MAXPARTS = 64;

__global__ void __launch_bounds__(256) part_count_kernel(const uint64_t *keys, size_t n, int nparts, unsigned long long *pcount)
{
    __shared__ unsigned int h[MAXPARTS];
    if (threadIdx.x < MAXPARTS) h[threadIdx.x] = 0;
    __syncthreads();
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) atomicAdd(&h[salsa_mix64(keys[i]) % (uint64_t)nparts], 1u);
    __syncthreads();
    if (threadIdx.x < nparts && h[threadIdx.x]) atomicAdd(&pcount[threadIdx.x], (unsigned long long)h[threadIdx.x]);
}

// cursor[p] starts at the exclusive prefix of pcount; each block reserves its share per part
__global__ void __launch_bounds__(256) part_scatter_kernel(const uint64_t *keys, const uint64_t *counts, size_t n, int nparts,
                                                          unsigned long long *cursor, uint64_t *okeys, uint64_t *ocounts)
{
    __shared__ unsigned int h[MAXPARTS];
    __shared__ unsigned long long base[MAXPARTS];
    size_t per = (n + gridDim.x - 1) / gridDim.x;
    size_t lo = (size_t)blockIdx.x * per, hi = lo + per < n ? lo + per : n;
    if (threadIdx.x < MAXPARTS) h[threadIdx.x] = 0;
    __syncthreads();
    for (size_t i = lo + threadIdx.x; i < hi; i += blockDim.x) atomicAdd(&h[salsa_mix64(keys[i]) % (uint64_t)nparts], 1u);
    __syncthreads();
    if (threadIdx.x < nparts) {
        base[threadIdx.x] = h[threadIdx.x] ? atomicAdd(&cursor[threadIdx.x], (unsigned long long)h[threadIdx.x]) : 0ull;
        h[threadIdx.x] = 0;
    }
    __syncthreads();
    for (size_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
        uint64_t k = keys[i];
        int p = (int)(salsa_mix64(k) % (uint64_t)nparts);
        unsigned long long o = base[p] + atomicAdd(&h[p], 1u);
        okeys[o] = k;
        ocounts[o] = counts[i];
    }
}

__global__ void sparse_add_kernel(const uint64_t *ikeys, const uint64_t *icounts, size_t n, uint64_t *keys, uint64_t *counts,
                                  uint64_t cap_mask, unsigned int *full)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t key = ikeys[i];
    uint64_t slot = salsa_mix64(key) & cap_mask;
    for (uint64_t probe = 0; probe <= cap_mask; probe++) {
        unsigned long long k = ((volatile unsigned long long *)keys)[slot];
        if (k == ~0ull) {
            unsigned long long old = atomicCAS((unsigned long long *)&keys[slot], ~0ull, (unsigned long long)key);
            k = (old == ~0ull || old == key) ? key : old;
        }
        if (k == key) {
            atomicAdd((unsigned long long *)&counts[slot], (unsigned long long)icounts[i]);
            return;
        }
        slot = (slot + 1) & cap_mask;
    }
    atomicExch(full, 1u);
}

}   // namespace

extern "C" int salsa_links_partition(const uint64_t *d_keys, const uint64_t *d_counts, size_t n, int32_t nparts,
                                     uint64_t *d_out_keys, uint64_t *d_out_counts, uint64_t *d_scratch, uint64_t *h_part_counts,
                                     void *stream)
{
    if (nparts < 1 || nparts > MAXPARTS) SALSA_FAIL(SALSA_ERR_ARG, "salsa_links_partition: 1..%d parts", MAXPARTS);
    if (!h_part_counts || !d_scratch || (n && (!d_keys || !d_counts || !d_out_keys || !d_out_counts)))
        SALSA_FAIL(SALSA_ERR_ARG, "salsa_links_partition: NULL argument");
    cudaStream_t st = as_stream(stream);
    unsigned long long *pc = (unsigned long long *)d_scratch;          // [nparts] counts, then cursors
    CUDA_TRY(cudaMemsetAsync(pc, 0, MAXPARTS * sizeof(uint64_t), st));
    unsigned grid = n ? (unsigned)((n + 256 * 16 - 1) / (256 * 16)) : 1;
    if (grid > 148 * 16) grid = 148 * 16;
    if (n) part_count_kernel<<<grid, 256, 0, st>>>(d_keys, n, nparts, pc);
    CUDA_TRY(cudaMemcpyAsync(h_part_counts, pc, (size_t)nparts * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    uint64_t cur[MAXPARTS], acc = 0;
    for (int p = 0; p < nparts; p++) { cur[p] = acc; acc += h_part_counts[p]; }
    CUDA_TRY(cudaMemcpyAsync(pc, cur, (size_t)nparts * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
    if (n) part_scatter_kernel<<<grid, 256, 0, st>>>(d_keys, d_counts, n, nparts, pc, d_out_keys, d_out_counts);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaStreamSynchronize(st));       // `cur` lives on this stack frame
    return SALSA_OK;
}

extern "C" int salsa_sparse_add(const uint64_t *d_in_keys, const uint64_t *d_in_counts, size_t n, uint64_t *d_keys,
                                uint64_t *d_counts, size_t capacity, uint32_t *d_full_flag, void *stream)
{
    if (n == 0) return SALSA_OK;
    if (!d_in_keys || !d_in_counts || !d_keys || !d_counts || !d_full_flag || capacity == 0 || (capacity & (capacity - 1)))
        SALSA_FAIL(SALSA_ERR_ARG, "salsa_sparse_add: bad argument");
    sparse_add_kernel<<<(unsigned)((n + 255) / 256), 256, 0, as_stream(stream)>>>(d_in_keys, d_in_counts, n, d_keys, d_counts,
                                                                                   capacity - 1, d_full_flag);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}

// ---- sparse merge, fused: hash table -> per-owner segments in ONE pass, no sort, no host round trip ----
// A segment is what one peer receives: [u64 count][u64 pad][u64 keys[slot_cap]][u32 counts[slot_cap]] -- 12 bytes per
// link, every segment the same size, so the exchange is ONE all-to-all with equal splits whose sizes the host knows
// without asking the device.  The receiver reads `count` from each segment it got.
namespace {

constexpr int PT = 256;          // threads per block
constexpr int PER = 8;           // table slots per thread and batch

__host__ __device__ inline size_t seg_bytes(size_t slot_cap) { return 16 + slot_cap * 12; }     // slot_cap % 4 == 0

__global__ void seg_init_kernel(uint8_t *seg, int nparts, size_t segb)
{
    if (threadIdx.x < nparts) {
        uint64_t *h = reinterpret_cast<uint64_t *>(seg + (size_t)threadIdx.x * segb);
        h[0] = 0;
        h[1] = 0;
    }
}

__global__ void __launch_bounds__(PT) seg_partition_kernel(const uint64_t *keys, const uint64_t *counts, size_t cap, int nparts,
                                                           size_t slot_cap, uint8_t *seg, size_t segb, unsigned int *overflow)
{
    __shared__ unsigned int h[MAXPARTS];
    __shared__ unsigned long long base[MAXPARTS];
    const size_t nbatch = (cap + (size_t)PT * PER - 1) / ((size_t)PT * PER);
    for (size_t b = blockIdx.x; b < nbatch; b += gridDim.x) {
        if (threadIdx.x < MAXPARTS) h[threadIdx.x] = 0;
        __syncthreads();
        uint64_t k[PER], c[PER];
        int part[PER];
        unsigned int rank[PER];
#pragma unroll
        for (int j = 0; j < PER; j++) {
            const size_t i = (b * PER + j) * PT + threadIdx.x;          // coalesced: consecutive threads, consecutive slots
            k[j] = i < cap ? keys[i] : ~0ull;
            c[j] = (k[j] != ~0ull) ? counts[i] : 0;
            part[j] = -1;
            if (c[j]) {
                part[j] = (int)(salsa_mix64(k[j]) % (uint64_t)nparts);
                rank[j] = atomicAdd(&h[part[j]], 1u);
            }
        }
        __syncthreads();
        if (threadIdx.x < nparts)
            base[threadIdx.x] = h[threadIdx.x]
                                    ? atomicAdd(reinterpret_cast<unsigned long long *>(seg + (size_t)threadIdx.x * segb),
                                                (unsigned long long)h[threadIdx.x])
                                    : 0ull;
        __syncthreads();
#pragma unroll
        for (int j = 0; j < PER; j++) {
            if (part[j] < 0) continue;
            const unsigned long long o = base[part[j]] + rank[j];
            if (o >= slot_cap || c[j] > 0xFFFFFFFFull) { atomicExch(overflow, 1u); continue; }
            uint8_t *sp = seg + (size_t)part[j] * segb + 16;
            reinterpret_cast<uint64_t *>(sp)[o] = k[j];
            reinterpret_cast<uint32_t *>(sp + slot_cap * 8)[o] = (uint32_t)c[j];
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) seg_add_kernel(const uint8_t *seg, int nparts, size_t slot_cap, size_t segb, uint64_t *keys,
                                                      uint64_t *counts, uint64_t cap_mask, unsigned int *full)
{
    const int p = blockIdx.y;
    const uint8_t *sp = seg + (size_t)p * segb;
    uint64_t n = *reinterpret_cast<const uint64_t *>(sp);
    if (n > slot_cap) n = slot_cap;                 // the sender flagged the overflow
    const uint64_t *ik = reinterpret_cast<const uint64_t *>(sp + 16);
    const uint32_t *ic = reinterpret_cast<const uint32_t *>(sp + 16 + slot_cap * 8);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const uint64_t key = ik[i];
        uint64_t slot = salsa_mix64(key) & cap_mask;
        bool done = false;
        for (uint64_t probe = 0; probe <= cap_mask; probe++) {
            unsigned long long k = ((volatile unsigned long long *)keys)[slot];
            if (k == ~0ull) {
                unsigned long long old = atomicCAS((unsigned long long *)&keys[slot], ~0ull, (unsigned long long)key);
                k = (old == ~0ull || old == key) ? key : old;
            }
            if (k == key) {
                atomicAdd((unsigned long long *)&counts[slot], (unsigned long long)ic[i]);
                done = true;
                break;
            }
            slot = (slot + 1) & cap_mask;
        }
        if (!done) atomicExch(full, 1u);
    }
}

}   // namespace

extern "C" size_t salsa_sparse_segment_bytes(size_t slot_cap) { return seg_bytes((slot_cap + 3) & ~(size_t)3); }

extern "C" int salsa_sparse_partition(const uint64_t *d_keys, const uint64_t *d_counts, size_t capacity, int32_t nparts,
                                      size_t slot_cap, void *d_segments, uint32_t *d_overflow, void *stream)
{
    if (nparts < 1 || nparts > MAXPARTS) SALSA_FAIL(SALSA_ERR_ARG, "salsa_sparse_partition: 1..%d parts", MAXPARTS);
    if (!d_keys || !d_counts || !d_segments || !d_overflow || capacity == 0 || slot_cap == 0 || (slot_cap & 3))
        SALSA_FAIL(SALSA_ERR_ARG, "salsa_sparse_partition: bad argument (slot_cap must be a multiple of 4)");
    cudaStream_t st = as_stream(stream);
    const size_t segb = seg_bytes(slot_cap);
    seg_init_kernel<<<1, MAXPARTS, 0, st>>>((uint8_t *)d_segments, nparts, segb);
    size_t nbatch = (capacity + (size_t)PT * PER - 1) / ((size_t)PT * PER);
    unsigned grid = (unsigned)(nbatch < 148 * 8 ? nbatch : 148 * 8);
    seg_partition_kernel<<<grid, PT, 0, st>>>(d_keys, d_counts, capacity, nparts, slot_cap, (uint8_t *)d_segments, segb, d_overflow);
    CUDA_TRY(cudaGetLastError());
    return SALSA_OK;
}

extern "C" int salsa_sparse_add_segments(const void *d_segments, int32_t nparts, size_t slot_cap, uint64_t *d_keys, uint64_t *d_counts,
                                         size_t capacity, uint32_t *d_full_flag, void *stream)
{
    if (nparts < 1 || nparts > MAXPARTS) SALSA_FAIL(SALSA_ERR_ARG, "salsa_sparse_add_segments: 1..%d parts", MAXPARTS);
    if (!d_segments || !d_keys || !d_counts || !d_full_flag || capacity == 0 || (capacity & (capacity - 1)) || slot_cap == 0 ||
        (slot_cap & 3))
        SALSA_FAIL(SALSA_ERR_ARG, "salsa_sparse_add_segments: bad argument");
    size_t per = (slot_cap + 255) / 256;
    dim3 grid((unsigned)(per < 148 * 4 ? per : 148 * 4), (unsigned)nparts);
    seg_add_kernel<<<grid, 256, 0, as_stream(stream)>>>((const uint8_t *)d_segments, nparts, slot_cap, seg_bytes(slot_cap), d_keys,
                                                        d_counts, capacity - 1, d_full_flag);
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
