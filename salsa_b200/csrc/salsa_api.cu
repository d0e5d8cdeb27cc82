// salsa_api.cu -- context, contig table, error plumbing of libsalsa_b200.
#include <stdarg.h>
#include <stdlib.h>

#include <algorithm>
#include <string>
#include <utility>
#include <vector>

#include "common.cuh"

typedef std::vector<std::pair<cudaEvent_t, cudaEvent_t> > EventPairs;      // salsa_timers

static thread_local char g_err[512] = "";

void salsa_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

extern "C" const char *salsa_last_error(void) { return g_err; }
extern "C" int salsa_version(void) { return 100; }

extern "C" int salsa_ctx_create(int device, salsa_ctx **out)
{
    if (!out) SALSA_FAIL(SALSA_ERR_ARG, "salsa_ctx_create: out is NULL");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        SALSA_FAIL(SALSA_ERR_CUDA, "no CUDA device available (%s); libsalsa_b200 has no CPU fallback",
                   e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    if (device < 0 || device >= ndev) SALSA_FAIL(SALSA_ERR_ARG, "device %d out of range (0..%d)", device, ndev - 1);
    CUDA_TRY(cudaSetDevice(device));
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        SALSA_FAIL(SALSA_ERR_UNSUPPORTED, "device %d is sm_%d%d; this library is built for sm_100a (B200) only", device,
                   prop.major, prop.minor);
    salsa_ctx *c = (salsa_ctx *)calloc(1, sizeof(salsa_ctx));
    if (!c) SALSA_FAIL(SALSA_ERR_NOMEM, "out of host memory");
    c->device = device;
    if (cudaMalloc(&c->d_mode, 16) != cudaSuccess || cudaMemset(c->d_mode, 0, 16) != cudaSuccess) {
        free(c);
        SALSA_FAIL(SALSA_ERR_CUDA, "salsa_ctx_create: cudaMalloc failed");
    }
    *out = c;
    return SALSA_OK;
}

static void free_table(salsa_ctx *c)
{
    free(c->h_blob);
    free(c->h_off);
    c->h_blob = NULL;
    c->h_off = NULL;
    cudaFree(c->d_slots);
    cudaFree(c->d_meta);
    cudaFree(c->d_blob);
    c->d_slots = c->d_meta = c->d_blob = NULL;
    free(c->h_meta);
    c->h_meta = NULL;
}

extern "C" void salsa_ctx_destroy(salsa_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    free_table(c);
    cudaFree(c->d_avoid);
    cudaFree(c->d_mode);
    delete (std::vector<std::string> *)c->h_involved;
    if (c->timer_events) {
        EventPairs *v = (EventPairs *)c->timer_events;
        for (size_t i = 0; i < v->size(); i++) { cudaEventDestroy((*v)[i].first); cudaEventDestroy((*v)[i].second); }
        delete v;
    }
    if (c->streams_ready) {
        cudaEventDestroy(c->ev_user);
        cudaStreamDestroy(c->s_copy);
        cudaStreamDestroy(c->s_comp);
        for (int i = 0; i < 2; i++) { cudaEventDestroy(c->ev_copy[i]); cudaEventDestroy(c->ev_comp[i]); }
    }
    free(c);
}

int salsa_ctx_streams(salsa_ctx *c)
{
    if (c->streams_ready) return SALSA_OK;
    CUDA_TRY(cudaStreamCreateWithFlags(&c->s_copy, cudaStreamNonBlocking));
    CUDA_TRY(cudaStreamCreateWithFlags(&c->s_comp, cudaStreamNonBlocking));
    for (int i = 0; i < 2; i++) {
        CUDA_TRY(cudaEventCreateWithFlags(&c->ev_copy[i], cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&c->ev_comp[i], cudaEventDisableTiming));
    }
    CUDA_TRY(cudaEventCreateWithFlags(&c->ev_user, cudaEventDisableTiming));
    c->streams_ready = true;
    return SALSA_OK;
}

extern "C" int32_t salsa_names_count(const salsa_ctx *c) { return c ? c->n : 0; }

// ---- timers (the reference has none; the driver's timings.tsv is per process) ----
int salsa_timer_begin(salsa_ctx *c, cudaStream_t st, void **slot)
{
    *slot = NULL;
    if (!c->timing) return SALSA_OK;
    if (!c->timer_events) c->timer_events = new EventPairs();
    cudaEvent_t a, b;
    CUDA_TRY(cudaEventCreate(&a));
    CUDA_TRY(cudaEventCreate(&b));
    CUDA_TRY(cudaEventRecord(a, st));
    EventPairs *v = (EventPairs *)c->timer_events;
    v->push_back(std::make_pair(a, b));
    *slot = (void *)(uintptr_t)v->size();          // index + 1
    return SALSA_OK;
}

void salsa_timer_end(salsa_ctx *c, cudaStream_t st, void *slot)
{
    if (!slot) return;
    EventPairs *v = (EventPairs *)c->timer_events;
    cudaEventRecord((*v)[(size_t)(uintptr_t)slot - 1].second, st);
}

extern "C" int salsa_timers_enable(salsa_ctx *c, int on)
{
    if (!c) SALSA_FAIL(SALSA_ERR_ARG, "salsa_timers_enable: ctx is NULL");
    c->timing = on ? 1 : 0;
    return SALSA_OK;
}

extern "C" int salsa_timers(salsa_ctx *c, double *out_ms, uint64_t *out_launches, int32_t n, int reset)
{
    if (!c || !out_ms || n < 1) SALSA_FAIL(SALSA_ERR_ARG, "salsa_timers: bad argument");
    CUDA_TRY(cudaSetDevice(c->device));
    EventPairs *v = (EventPairs *)c->timer_events;
    if (v) {
        for (size_t i = 0; i < v->size(); i++) {
            CUDA_TRY(cudaEventSynchronize((*v)[i].second));
            float ms = 0;
            CUDA_TRY(cudaEventElapsedTime(&ms, (*v)[i].first, (*v)[i].second));
            c->timer_ms[0] += ms;
            c->timer_n[0]++;
            cudaEventDestroy((*v)[i].first);
            cudaEventDestroy((*v)[i].second);
        }
        v->clear();
    }
    for (int32_t i = 0; i < n; i++) {
        out_ms[i] = i < 4 ? c->timer_ms[i] : 0.0;
        if (out_launches) out_launches[i] = i < 4 ? c->timer_n[i] : 0;
    }
    if (reset)
        for (int i = 0; i < 4; i++) { c->timer_ms[i] = 0; c->timer_n[i] = 0; }
    return SALSA_OK;
}

extern "C" int salsa_ctx_scan_mode(salsa_ctx *c, int mode)
{
    if (!c || mode < 0 || mode > 2) SALSA_FAIL(SALSA_ERR_ARG, "salsa_ctx_scan_mode: mode must be 0 (auto), 1 (exact) or 2 (cheap)");
    c->scan_mode = mode;
    return SALSA_OK;
}

extern "C" int salsa_names_build(salsa_ctx *c, const char *blob, const int64_t *off, int32_t n, const int64_t *len_floor,
                                 const uint32_t *flags)
{
    if (!c || !blob || !off || n < 0 || (n > 0 && (!len_floor || !flags))) SALSA_FAIL(SALSA_ERR_ARG, "salsa_names_build: bad argument");
    CUDA_TRY(cudaSetDevice(c->device));
    // ids must reproduce Python-2 byte-string '<' (make_links.py:90): verify strict order
    for (int32_t i = 0; i < n; i++) {
        int64_t li = off[i + 1] - off[i];
        if (li <= 0 || li >= SALSA_MAX_LINE) SALSA_FAIL(SALSA_ERR_ARG, "contig name %d has length %lld", i, (long long)li);
        if (i > 0) {
            int64_t lp = off[i] - off[i - 1];
            int cmp = memcmp(blob + off[i - 1], blob + off[i], (size_t)std::min(li, lp));
            if (cmp > 0 || (cmp == 0 && lp >= li))
                SALSA_FAIL(SALSA_ERR_ARG, "contig names must be strictly increasing in byte order (entry %d)", i);
        }
    }
    if (off[n] >= (int64_t)1 << 32) SALSA_FAIL(SALSA_ERR_ARG, "name blob too large");
    free_table(c);
    uint32_t nslots = 64;
    // Sparse on purpose: a second probe of one lane is a second L2 round trip (and a divergent loop) for the whole warp of
    // the scanner's phase 2 (bed_stream.cuh), and with 32 lookups per batch "some lane collides" is the rule at the usual
    // load of 0.5 (measured on B200, 3,000 contigs: 22.0 ms per 47 GB step at load 0.37, 21.4 at 0.09, 21.0 at 0.01).
    // 64 slots per name up to 2^20 slots (32 MB, stays in L2); beyond that at least 2 per name.
    while ((nslots < (uint32_t)n * 64u + 2u && nslots < (1u << 20)) || nslots < (uint32_t)n * 2u + 2u) nslots <<= 1;
    std::vector<NameSlot> slots(nslots);
    for (auto &s : slots) {
        memset(&s, 0, sizeof s);
        s.len = 0xFFFFFFFFu;
        s.id = -1;
    }
    for (int32_t i = 0; i < n; i++) {
        uint32_t len = (uint32_t)(off[i + 1] - off[i]);
        uint32_t w[4] = {0, 0, 0, 0};
        memcpy(w, blob + off[i], std::min<uint32_t>(len, 16));
        uint32_t h = salsa_name_hash(w[0], w[1], w[2], w[3], len) & (nslots - 1);
        while (slots[h].len != 0xFFFFFFFFu) h = (h + 1) & (nslots - 1);
        memcpy(slots[h].w, w, 16);
        slots[h].len = len;
        slots[h].id = i;
        slots[h].blob_off = (uint32_t)off[i];
    }
    c->h_meta = (ContigMeta *)calloc((size_t)n + 1, sizeof(ContigMeta));
    if (!c->h_meta) SALSA_FAIL(SALSA_ERR_NOMEM, "out of host memory");
    for (int32_t i = 0; i < n; i++) {
        c->h_meta[i].len_floor = len_floor[i];
        c->h_meta[i].flags = flags[i] & (SALSA_CTG_HAS_LEN | SALSA_CTG_HAS_RE);
    }
    CUDA_TRY(cudaMalloc(&c->d_slots, nslots * sizeof(NameSlot)));
    CUDA_TRY(cudaMalloc(&c->d_meta, ((size_t)n + 1) * sizeof(ContigMeta)));
    CUDA_TRY(cudaMalloc(&c->d_blob, (size_t)off[n] + 16));
    CUDA_TRY(cudaMemcpy(c->d_slots, slots.data(), nslots * sizeof(NameSlot), cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(c->d_meta, c->h_meta, ((size_t)n + 1) * sizeof(ContigMeta), cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(c->d_blob, blob, (size_t)off[n], cudaMemcpyHostToDevice));
    c->h_blob = (char *)malloc((size_t)off[n] + 1);
    c->h_off = (int64_t *)malloc(((size_t)n + 1) * sizeof(int64_t));
    if (!c->h_blob || !c->h_off) SALSA_FAIL(SALSA_ERR_NOMEM, "out of host memory");
    memcpy(c->h_blob, blob, (size_t)off[n]);
    memcpy(c->h_off, off, ((size_t)n + 1) * sizeof(int64_t));
    delete (std::vector<std::string> *)c->h_involved;
    c->h_involved = NULL;
    c->n = n;
    c->names_punct = 0;
    for (int64_t i = 0; i < off[n]; i++)
        if ((unsigned char)blob[i] - 0x21u <= 0x0Du) { c->names_punct = 1; break; }
    c->t.slots = (const NameSlot *)c->d_slots;
    c->t.mask = nslots - 1;
    c->t.meta = (const ContigMeta *)c->d_meta;
    c->t.blob = (const char *)c->d_blob;
    c->t.n = n;
    c->t.avoid = NULL;
    c->t.navoid = 0;
    cudaFree(c->d_avoid);
    c->d_avoid = NULL;
    return SALSA_OK;
}

extern "C" int salsa_avoid_set(salsa_ctx *c, const int32_t *pairs, int32_t npairs)
{
    if (!c || npairs < 0 || (npairs > 0 && !pairs)) SALSA_FAIL(SALSA_ERR_ARG, "salsa_avoid_set: bad argument");
    CUDA_TRY(cudaSetDevice(c->device));
    cudaFree(c->d_avoid);
    c->d_avoid = NULL;
    c->t.avoid = NULL;
    c->t.navoid = 0;
    for (int32_t i = 0; i < c->n; i++) c->h_meta[i].flags &= ~CTG_INVOLVED;
    std::vector<uint64_t> keys;
    for (int32_t i = 0; i < npairs; i++) {
        int32_t a = pairs[2 * i], b = pairs[2 * i + 1];
        if (a < 0 || a >= c->n || b < 0 || b >= c->n) SALSA_FAIL(SALSA_ERR_ARG, "avoid pair %d: id out of range", i);
        if (a == b) continue;   // same contig: make_links.py:71 never consults the set
        keys.push_back(((uint64_t)(uint32_t)a << 32) | (uint32_t)b);
        keys.push_back(((uint64_t)(uint32_t)b << 32) | (uint32_t)a);
        c->h_meta[a].flags |= CTG_INVOLVED;
        c->h_meta[b].flags |= CTG_INVOLVED;
    }
    std::sort(keys.begin(), keys.end());
    keys.erase(std::unique(keys.begin(), keys.end()), keys.end());
    {
        std::vector<std::string> *inv = new std::vector<std::string>();
        for (int32_t i = 0; i < c->n; i++)
            if (c->h_meta[i].flags & CTG_INVOLVED) inv->push_back(std::string(c->h_blob + c->h_off[i], (size_t)(c->h_off[i + 1] - c->h_off[i])));
        std::sort(inv->begin(), inv->end());
        delete (std::vector<std::string> *)c->h_involved;
        c->h_involved = inv;
    }
    if (c->n > 0)
        CUDA_TRY(cudaMemcpy(c->d_meta, c->h_meta, ((size_t)c->n + 1) * sizeof(ContigMeta), cudaMemcpyHostToDevice));
    if (!keys.empty()) {
        CUDA_TRY(cudaMalloc(&c->d_avoid, keys.size() * sizeof(uint64_t)));
        CUDA_TRY(cudaMemcpy(c->d_avoid, keys.data(), keys.size() * sizeof(uint64_t), cudaMemcpyHostToDevice));
        c->t.avoid = (const uint64_t *)c->d_avoid;
        c->t.navoid = (int32_t)keys.size();
    }
    return SALSA_OK;
}
