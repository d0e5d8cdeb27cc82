// ubench_pipes.cu -- issue-rate micro-benchmark of the integer / shuffle / shared-memory instructions the
// BED scanner is made of (sm_100a).  Every test runs 8 independent dependency chains per thread with
// 8 warps per SM sub-partition (4 CTAs of 256 threads per SM) and reports warp instructions per clock per
// sub-partition (1.0 = the issue limit), from the difference of two run lengths (launch overhead cancels).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_pipes tools/ubench_pipes.cu
// Results on B200: profiles/r02_ubench_pipes.jsonl
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define CHAIN8(STMT) STMT(0) STMT(1) STMT(2) STMT(3) STMT(4) STMT(5) STMT(6) STMT(7)
#define ALT16(S, T) S(0) T(1) S(2) T(3) S(4) T(5) S(6) T(7) S(1) T(0) S(3) T(2) S(5) T(4) S(7) T(6)

#define I_LOP3(j) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[j]) : "r"(b), "r"(c));
#define I_ADD(j) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[j]) : "r"(b));
#define I_IMAD(j) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[j]) : "r"(b), "r"(c));
#define I_SHF(j) asm volatile("shf.r.wrap.b32 %0, %0, %1, %2;" : "+r"(a[j]) : "r"(b), "r"(c));
#define I_PRMT(j) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(a[j]) : "r"(b), "r"(c));
#define I_DP4A(j) asm volatile("dp4a.u32.u32 %0, %0, %1, %2;" : "+r"(a[j]) : "r"(b), "r"(c));
#define I_POPC(j) asm volatile("popc.b32 %0, %0;" : "+r"(a[j]));
#define I_CLZ(j) asm volatile("clz.b32 %0, %0;" : "+r"(a[j]));
#define I_BREV(j) asm volatile("brev.b32 %0, %0;" : "+r"(a[j]));
#define I_BALLOT(j) asm volatile("{.reg .pred p; setp.ne.u32 p, %0, 0; vote.sync.ballot.b32 %0, p, 0xffffffff;}" : "+r"(a[j]));
#define I_SHFL(j) asm volatile("shfl.sync.up.b32 %0, %0, 1, 0, 0xffffffff;" : "+r"(a[j]));
#define I_LDS(j) asm volatile("{.reg .u32 t; and.b32 t, %0, 0x3ffc; add.u32 t, t, %1; ld.shared.u32 %0, [t];}" : "+r"(a[j]) : "r"(saddr));
#define I_LDS128(j) asm volatile("{.reg .u32 t, x, y, z; and.b32 t, %0, 0x3ff0; add.u32 t, t, %1; ld.shared.v4.u32 {%0, x, y, z}, [t];}" : "+r"(a[j]) : "r"(saddr));
#define I_VABS(j) asm volatile("vabsdiff4.u32.u32.u32 %0, %0, %1, %2;" : "+r"(a[j]) : "r"(b), "r"(c));
#define I_SETSEL(j) asm volatile("{.reg .pred p; setp.lt.u32 p, %0, %1; selp.u32 %0, %2, %0, p;}" : "+r"(a[j]) : "r"(b), "r"(c));

template <int OP>
__global__ void __launch_bounds__(256) k(uint32_t *out, uint32_t seed, int iters)
{
    __shared__ uint32_t sm[4096];
    for (int i = threadIdx.x; i < 4096; i += 256) sm[i] = i * 2654435761u + seed;
    __syncthreads();
    uint32_t a[8], b = seed | 1u, c = seed * 3u + 7u;
#pragma unroll
    for (int j = 0; j < 8; j++) a[j] = threadIdx.x * 17u + j * 1315423911u + seed;
    uint32_t saddr = (uint32_t)__cvta_generic_to_shared(sm);
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
        if (OP == 0) { CHAIN8(I_LOP3) CHAIN8(I_LOP3) }
        else if (OP == 1) { CHAIN8(I_ADD) CHAIN8(I_ADD) }          // ptxas folds two adds into one IADD3
        else if (OP == 2) { CHAIN8(I_IMAD) CHAIN8(I_IMAD) }
        else if (OP == 3) { CHAIN8(I_SHF) CHAIN8(I_SHF) }
        else if (OP == 4) { CHAIN8(I_PRMT) CHAIN8(I_PRMT) }
        else if (OP == 5) { CHAIN8(I_DP4A) CHAIN8(I_DP4A) }
        else if (OP == 6) { CHAIN8(I_POPC) CHAIN8(I_POPC) }
        else if (OP == 7) { CHAIN8(I_CLZ) CHAIN8(I_CLZ) }
        else if (OP == 8) { CHAIN8(I_BREV) CHAIN8(I_BREV) }
        else if (OP == 9) { CHAIN8(I_BALLOT) CHAIN8(I_BALLOT) }
        else if (OP == 10) { CHAIN8(I_SHFL) CHAIN8(I_SHFL) }
        else if (OP == 11) { CHAIN8(I_LDS) CHAIN8(I_LDS) }
        else if (OP == 12) { CHAIN8(I_VABS) CHAIN8(I_VABS) }
        else if (OP == 13) { CHAIN8(I_SETSEL) CHAIN8(I_SETSEL) }
        else if (OP == 14) { ALT16(I_LOP3, I_IMAD) }
        else if (OP == 15) { ALT16(I_LOP3, I_DP4A) }
        else if (OP == 16) { ALT16(I_LOP3, I_LDS) }
        else if (OP == 17) { ALT16(I_SHF, I_IMAD) }
        else if (OP == 18) { ALT16(I_DP4A, I_IMAD) }
        else if (OP == 19) { CHAIN8(I_LDS128) CHAIN8(I_LDS128) }
        else if (OP == 20) { ALT16(I_LOP3, I_POPC) }
        else if (OP == 21) { ALT16(I_LOP3, I_SHFL) }
        else if (OP == 22) { ALT16(I_IMAD, I_SHFL) }
    }
    uint32_t r = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) r ^= a[j];
    out[blockIdx.x * 256 + threadIdx.x] = r;
}

template <int OP>
float once(int iters, uint32_t *out)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0);
        k<OP><<<148 * 4, 256>>>(out, 12345u, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (r > 0 && ms < best) best = ms;
    }
    return best;
}

template <int OP>
void run(const char *name, double instr_per_slot, uint32_t *out, double clk)
{
    float a = once<OP>(2048, out), b = once<OP>(10240, out);
    double per_smsp = 8.0 * 8192.0 * 16.0 * instr_per_slot / ((b - a) * 1e-3 * clk);
    printf("{\"op\": \"%s\", \"warp_instr_per_clk_per_smsp\": %.3f, \"ms_2048\": %.4f, \"ms_10240\": %.4f, \"err\": \"%s\"}\n", name,
           per_smsp, a, b, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    uint32_t *out;
    cudaMalloc(&out, 148 * 4 * 256 * 4);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    double clk = khz * 1e3;
    run<0>("lop3", 1, out, clk);
    run<1>("iadd3 (two adds each)", 0.5, out, clk);
    run<2>("imad", 1, out, clk);
    run<3>("shf", 1, out, clk);
    run<4>("prmt", 1, out, clk);
    run<5>("dp4a", 1, out, clk);
    run<6>("popc", 1, out, clk);
    run<7>("clz (flo)", 1, out, clk);
    run<8>("brev", 1, out, clk);
    run<9>("setp+ballot", 2, out, clk);
    run<10>("shfl.up", 1, out, clk);
    run<11>("and+add+lds32 (random banks)", 3, out, clk);
    run<12>("vabsdiff4", 1, out, clk);
    run<13>("setp+selp", 2, out, clk);
    run<14>("lop3|imad", 1, out, clk);
    run<15>("lop3|dp4a", 1, out, clk);
    run<16>("lop3|(and+add+lds32)", 2, out, clk);
    run<17>("shf|imad", 1, out, clk);
    run<18>("dp4a|imad", 1, out, clk);
    run<19>("and+add+lds128 (random banks)", 3, out, clk);
    run<20>("lop3|popc", 1, out, clk);
    run<21>("lop3|shfl", 1, out, clk);
    run<22>("imad|shfl", 1, out, clk);
    return 0;
}
