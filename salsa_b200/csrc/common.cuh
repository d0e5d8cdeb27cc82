// common.cuh -- shared declarations of libsalsa_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/salsa_b200.h"

// ---------------------------------------------------------------- errors
void salsa_set_error(const char *fmt, ...);

#define SALSA_FAIL(code, ...)          \
    do {                               \
        salsa_set_error(__VA_ARGS__);  \
        return (code);                 \
    } while (0)

#define CUDA_TRY(expr)                                                                     \
    do {                                                                                   \
        cudaError_t e__ = (expr);                                                          \
        if (e__ != cudaSuccess)                                                            \
            SALSA_FAIL(SALSA_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), \
                       __FILE__, __LINE__);                                                \
    } while (0)

// ---------------------------------------------------------------- contig table
// One slot of the device name table (open addressing, linear probing).
// Names up to 16 bytes are matched entirely from the slot; longer names are
// verified against the blob.
struct __align__(16) NameSlot {
    uint32_t w[4];      // first 16 name bytes, zero padded, little endian
    uint32_t len;       // name length; 0xFFFFFFFF = empty slot
    int32_t id;
    uint32_t blob_off;  // offset of the name in the blob (for len > 16)
    uint32_t pad;
};

struct ContigMeta {
    int64_t len_floor;
    uint32_t flags;     // SALSA_CTG_* | CTG_INVOLVED
    uint32_t pad;
};
#define CTG_INVOLVED 0x100u   // contig occurs in some avoid pair

struct DevTable {
    const NameSlot *slots;
    uint32_t mask;            // nslots - 1
    const ContigMeta *meta;
    const char *blob;
    int32_t n;
    // avoid set: sorted uint64 (a << 32 | b), both orders present
    const uint64_t *avoid;
    int32_t navoid;
};

struct salsa_ctx {
    int device;
    DevTable t;
    void *d_slots, *d_meta, *d_blob, *d_avoid;
    ContigMeta *h_meta;       // host copy (avoid-set rebuilds update flags)
    int32_t n;
    // lazily created by the host-streaming entry points
    cudaStream_t s_copy, s_comp;
    cudaEvent_t ev_copy[2], ev_comp[2];
    bool streams_ready;
    // streaming scanner: special-byte test (0 auto, 1 exact, 2 cheap), contig names with a byte in 0x21..0x2E, device flag
    int scan_mode, names_punct;
    int *d_mode;
    // host copy of the names of contigs that occur in an avoid pair, sorted (salsa_bed_ingest_dense_host cuts its
    // chunks at lines that can never be skipped); '\0'-separated blob + offsets
    char *h_blob;
    int64_t *h_off;
    void *h_involved;          // std::vector<std::string> *
    cudaEvent_t ev_user;       // orders the internal streams after the caller's stream
    // salsa_timers: when enabled, every BED scan launch is bracketed by a pair of events
    int timing;
    void *timer_events;        // std::vector<std::pair<cudaEvent_t, cudaEvent_t>> *
    double timer_ms[4];        // accumulated: [0] BED scan kernels
    uint64_t timer_n[4];
};
int salsa_timer_begin(salsa_ctx *c, cudaStream_t st, void **slot);
void salsa_timer_end(salsa_ctx *c, cudaStream_t st, void *slot);
int salsa_ctx_streams(salsa_ctx *c);

__host__ __device__ inline uint32_t salsa_rotl(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }

__host__ __device__ inline uint32_t salsa_name_hash(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t len)
{
    uint32_t h = (w0 * 0x9E3779B1u) ^ salsa_rotl(w1 * 0x85EBCA77u, 13) ^ (w2 * 0xC2B2AE3Du) ^
                 salsa_rotl(w3 * 0x27D4EB2Fu, 17) ^ (len * 0x165667B1u);
    h ^= h >> 15;
    h *= 0x2C1B3C6Du;
    h ^= h >> 13;
    return h;
}

__host__ __device__ inline uint64_t salsa_mix64(uint64_t x)
{
    x ^= x >> 33;
    x *= 0xff51afd7ed558ccdull;
    x ^= x >> 33;
    x *= 0xc4ceb9fe1a85ec53ull;
    x ^= x >> 33;
    return x;
}

static inline cudaStream_t as_stream(void *s) { return (cudaStream_t)s; }
