// Microbenchmark: issue throughput of the packed-integer / DPX / fp16x2 instructions the DP kernels are built from,
// alone and in pairs, in warp-instructions per clock per SM sub-partition.  Build: nvcc -arch=sm_100a -O3 pipes.cu
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#define CHAINS 8
#define ITERS 4096

template <int OP>
__device__ __forceinline__ void step(uint32_t (&x)[CHAINS], uint32_t (&y)[CHAINS], uint32_t c, uint32_t d)
{
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) {
        if (OP == 0) x[k] = __vmaxs2(x[k], c);                              // VIMNMX.S16x2
        if (OP == 1) x[k] = __viaddmax_s16x2(x[k], c, d);                   // VIADDMNMX.S16x2
        if (OP == 2) x[k] = __vimax3_s16x2(x[k], c, d);                     // VIMNMX3.S16x2
        if (OP == 3) x[k] = __vadd2(x[k], c);                               // VIADD.16x2
        if (OP == 4) asm volatile("add.u32 %0, %0, %1;" : "+r"(x[k]) : "r"(c));          // IADD3 or IMAD.IADD (ptxas decides)
        if (OP == 5) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[k]) : "r"(c), "r"(d));   // IMAD
        if (OP == 6) x[k] = __byte_perm(x[k], c, d);                        // PRMT
        if (OP == 7) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[k]) : "r"(c), "r"(d)); // LOP3
        if (OP == 8) { __half2 h = *reinterpret_cast<__half2 *>(&x[k]); h = __hmax2(h, *reinterpret_cast<__half2 *>(&c)); x[k] = *reinterpret_cast<uint32_t *>(&h); }   // HMNMX2
        if (OP == 9) { __half2 h = *reinterpret_cast<__half2 *>(&x[k]); h = __hadd2(h, *reinterpret_cast<__half2 *>(&c)); x[k] = *reinterpret_cast<uint32_t *>(&h); }   // HADD2
        if (OP == 10) { x[k] = __viaddmax_s16x2(x[k], c, d); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(y[k]) : "r"(c), "r"(d)); }   // DPX + IMAD
        if (OP == 11) { x[k] = __viaddmax_s16x2(x[k], c, d); __half2 h = *reinterpret_cast<__half2 *>(&y[k]); h = __hmax2(h, *reinterpret_cast<__half2 *>(&c)); y[k] = *reinterpret_cast<uint32_t *>(&h); }   // DPX + HMNMX2
        if (OP == 12) { x[k] = __viaddmax_s16x2(x[k], c, d); y[k] = __byte_perm(y[k], c, d); }   // DPX + PRMT
        if (OP == 13) { x[k] = __viaddmax_s16x2(x[k], c, d); __half2 h = *reinterpret_cast<__half2 *>(&y[k]); h = __hadd2(h, *reinterpret_cast<__half2 *>(&c)); y[k] = *reinterpret_cast<uint32_t *>(&h); }   // DPX + HADD2
        if (OP == 14) { bool p, q; x[k] = __vibmax_s16x2(x[k], c, &p, &q); if (p) y[k] += 1; }   // VIMNMX with predicate out + predicated add
        if (OP == 15) asm volatile("sub.u32 %0, %0, %1;" : "+r"(x[k]) : "r"(c));
        if (OP == 16) { x[k] = __vmaxs2(x[k], c); asm volatile("add.u32 %0, %0, %1;" : "+r"(y[k]) : "r"(d)); }
        if (OP == 17) { __half2 h = *reinterpret_cast<__half2 *>(&x[k]); h = __hfma2(h, *reinterpret_cast<__half2 *>(&c), *reinterpret_cast<__half2 *>(&d)); x[k] = *reinterpret_cast<uint32_t *>(&h); }   // HFMA2
        if (OP == 18) { x[k] = __vmaxs2(x[k], c); __half2 h = *reinterpret_cast<__half2 *>(&y[k]); h = __hfma2(h, *reinterpret_cast<__half2 *>(&c), *reinterpret_cast<__half2 *>(&d)); y[k] = *reinterpret_cast<uint32_t *>(&h); }
    }
}

template <int OP>
__device__ __forceinline__ void step2(uint32_t (&x)[CHAINS], uint32_t (&y)[CHAINS], uint32_t c, uint32_t d, uint32_t c2, uint32_t d2)
{
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) {
        if (OP == 10) { x[k] = __viaddmax_s16x2(x[k], c, d); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(y[k]) : "r"(c2), "r"(d2)); }
        if (OP == 11) { x[k] = __viaddmax_s16x2(x[k], c, d); __half2 h = *reinterpret_cast<__half2 *>(&y[k]); h = __hmax2(h, *reinterpret_cast<__half2 *>(&c2)); y[k] = *reinterpret_cast<uint32_t *>(&h); }
        if (OP == 12) { x[k] = __viaddmax_s16x2(x[k], c, d); y[k] = __byte_perm(y[k], c2, d2); }
        if (OP == 13) { x[k] = __viaddmax_s16x2(x[k], c, d); __half2 h = *reinterpret_cast<__half2 *>(&y[k]); h = __hadd2(h, *reinterpret_cast<__half2 *>(&c2)); y[k] = *reinterpret_cast<uint32_t *>(&h); }
        if (OP == 16) { x[k] = __vmaxs2(x[k], c); asm volatile("add.u32 %0, %0, %1;" : "+r"(y[k]) : "r"(d2)); }
        if (OP == 18) { x[k] = __vmaxs2(x[k], c); __half2 h = *reinterpret_cast<__half2 *>(&y[k]); h = __hfma2(h, *reinterpret_cast<__half2 *>(&c2), *reinterpret_cast<__half2 *>(&d2)); y[k] = *reinterpret_cast<uint32_t *>(&h); }
        if (OP == 19) { x[k] = __vmaxs2(x[k], c); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(y[k]) : "r"(c2), "r"(d2)); }
        if (OP == 20) { x[k] = __vmaxs2(x[k], c); y[k] = __vmaxs2(y[k], c2); }
        if (OP == 21) { x[k] = __vmaxs2(x[k], c); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(y[k]) : "r"(c2), "n"(12345)); }
        if (OP == 22) { x[k] = __vmaxs2(x[k], c); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(y[k]) : "r"(c2), "r"(d2)); }
        if (OP == 23) { x[k] = __viaddmax_s16x2(x[k], c, d); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(y[k]) : "r"(c2), "n"(12345)); }
        if (OP == 24) { x[k] = __viaddmax_s16x2(x[k], c, d); asm volatile("add.u32 %0, %0, %1;" : "+r"(y[k]) : "r"(d2)); }
    }
}

template <int OP>
__global__ void __launch_bounds__(1024) bench(uint32_t *out, uint32_t c, uint32_t d, long long *clk)
{
    uint32_t x[CHAINS], y[CHAINS];
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) { x[k] = threadIdx.x * 7 + k; y[k] = threadIdx.x + 3 * k; }
    uint32_t c2 = c + threadIdx.x, d2 = d ^ threadIdx.x;
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) { if (OP >= 10 && OP != 14) step2<OP>(x, y, c, d, c2, d2); else step<OP>(x, y, c, d); }
    long long t1 = clock64();
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) s ^= x[k] ^ y[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char *name, int n_instr_per_step, uint32_t *out, long long *clk, int warps_per_smsp)
{
    const int threads = 32 * 4 * warps_per_smsp > 1024 ? 1024 : 32 * 4 * warps_per_smsp;
    const int blocks_per_sm = (32 * 4 * warps_per_smsp) / threads;
    const int blocks = 148 * blocks_per_sm;
    bench<OP><<<blocks, threads>>>(out, 0x00030005u, 0x00010002u, clk);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    bench<OP><<<blocks, threads>>>(out, 0x00030005u, 0x00010002u, clk);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long h[8]; cudaMemcpy(h, clk, sizeof h, cudaMemcpyDeviceToHost);
    const double instr_per_warp = (double)ITERS * CHAINS * n_instr_per_step;
    // per SMSP: warps_per_smsp warps each issue instr_per_warp instructions in h[0] clocks
    printf("%-28s warps/SMSP=%d  clk/warp-instr/SMSP = %.3f   (%.3f ms, err=%s)\n", name, warps_per_smsp,
           (double)h[0] / (instr_per_warp * warps_per_smsp), ms, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    uint32_t *out; long long *clk;
    cudaMalloc(&out, 148 * 4 * 1024 * 4); cudaMalloc(&clk, 148 * 8 * 8);
    for (int w : {2, 4, 8}) {
        run<0>("VIMNMX.S16x2", 1, out, clk, w);
        run<1>("VIADDMNMX.S16x2", 1, out, clk, w);
        run<2>("VIMNMX3.S16x2", 1, out, clk, w);
        run<3>("VIADD.16x2", 1, out, clk, w);
        run<4>("add.u32", 1, out, clk, w);
        run<15>("sub.u32", 1, out, clk, w);
        run<5>("IMAD", 1, out, clk, w);
        run<6>("PRMT", 1, out, clk, w);
        run<7>("LOP3", 1, out, clk, w);
        run<8>("HMNMX2", 1, out, clk, w);
        run<9>("HADD2", 1, out, clk, w);
        run<17>("HFMA2", 1, out, clk, w);
        run<10>("VIADDMNMX + IMAD", 2, out, clk, w);
        run<11>("VIADDMNMX + HMNMX2", 2, out, clk, w);
        run<12>("VIADDMNMX + PRMT", 2, out, clk, w);
        run<13>("VIADDMNMX + HADD2", 2, out, clk, w);
        run<16>("VIMNMX + add.u32", 2, out, clk, w);
        run<18>("VIMNMX + HFMA2", 2, out, clk, w);
        run<14>("VIMNMX.pred + @P add", 2, out, clk, w);
        run<19>("VIMNMX + IMAD(rrr)", 2, out, clk, w);
        run<20>("VIMNMX + VIMNMX", 2, out, clk, w);
        run<21>("VIMNMX + IMAD(rr,imm)", 2, out, clk, w);
        run<22>("VIMNMX + LOP3", 2, out, clk, w);
        run<23>("VIADDMNMX + IMAD(rr,imm)", 2, out, clk, w);
        run<24>("VIADDMNMX + add.u32", 2, out, clk, w);
    }
    return 0;
}
