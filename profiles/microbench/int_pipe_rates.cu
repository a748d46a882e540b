// Integer-pipe rate microbenchmark for sm_100a: how many SM cycles one warp
// instruction of each multiply flavour occupies.  Decides how the 64-bit Shoup
// multiply of the NTT butterfly is decomposed (DESIGN.md "integer ceiling").
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o int_pipe_rates int_pipe_rates.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 256
#define CHAINS 8

template <int OP>
__global__ void __launch_bounds__(1024) k(uint32_t* out, uint32_t seed, long long* cyc) {
    extern __shared__ uint32_t dummy_smem[];
    uint32_t a[CHAINS], b[CHAINS];
    uint64_t w[CHAINS];
    double d[CHAINS];
#pragma unroll
    for (int i = 0; i < CHAINS; i++) {
        a[i] = seed * (threadIdx.x + 1) + i;
        b[i] = seed + 77 * i + threadIdx.x;
        w[i] = ((uint64_t)a[i] << 32) | b[i];
        d[i] = (double)a[i];
    }
    long long t0 = clock64();
#pragma unroll 4
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < CHAINS; i++) {
            if (OP == 0) asm volatile("mad.lo.u32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(b[i]));
            if (OP == 1) asm volatile("mad.hi.u32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(b[i]));
            if (OP == 2) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(a[i]), "r"(b[i]));
            if (OP == 3) asm volatile("mul.hi.u64 %0, %0, %1;" : "+l"(w[i]) : "l"(w[(i + 1) % CHAINS] | 1));
            if (OP == 4) asm volatile("mad.lo.u64 %0, %0, %1, %0;" : "+l"(w[i]) : "l"(w[(i + 1) % CHAINS] | 1));
            if (OP == 5) asm volatile("add.cc.u32 %0, %0, %1;\n\taddc.u32 %1, %1, %0;" : "+r"(a[i]), "+r"(b[i]));
            if (OP == 6) asm volatile("fma.rn.f64 %0, %0, %1, %0;" : "+d"(d[i]) : "d"(1.0000001));
            if (OP == 7) {  // mixed: one wide + one add pair (do the pipes overlap?)
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(a[i]), "r"(b[i]));
                asm volatile("add.cc.u32 %0, %0, %1;\n\taddc.u32 %1, %1, %0;" : "+r"(a[i]), "+r"(b[i]));
            }
            if (OP == 8) asm volatile("mul.wide.u16 %0, %1, %2;" : "=r"(a[i]) : "h"((uint16_t)a[i]), "h"((uint16_t)b[i]));
            if (OP == 9) {  // lo + hi pair instead of one wide
                uint32_t lo, hi;
                asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(lo) : "r"(a[i]), "r"(b[i]));
                asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(hi) : "r"(a[i]), "r"(b[i]));
                a[i] = lo ^ hi;
            }
        }
    }
    long long t1 = clock64();
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; i++) acc ^= a[i] ^ b[i] ^ (uint32_t)w[i] ^ (uint32_t)(w[i] >> 32) ^ (uint32_t)d[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) atomicMax((unsigned long long*)cyc, (unsigned long long)(t1 - t0));
}

template <int OP>
void run(const char* name, int per_iter) {
    uint32_t* out;
    long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4);
    cudaMallocManaged(&cyc, 8);
    cudaFuncSetAttribute(k<OP>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024);
    for (int warps = 8; warps <= 32; warps *= 2) {   // 1 CTA per SM (120 KB smem), CTA = warps*32 threads
        for (int rep = 0; rep < 3; rep++) {
            *cyc = 0;
            k<OP><<<148, warps * 32, 120 * 1024>>>(out, 12345, cyc);
            cudaDeviceSynchronize();
        }
        double warp_instr_per_smsp = (double)ITERS * CHAINS * per_iter * warps / 4.0;
        printf("%-28s warps/SM=%2d  cycles=%8lld  cyc/warp-instr/SMSP=%.2f\n", name, warps, *cyc, *cyc / warp_instr_per_smsp);
    }
    cudaFree(out);
    cudaFree(cyc);
}

int main() {
    run<0>("mad.lo.u32 (IMAD)", 1);
    run<1>("mad.hi.u32 (IMAD.HI)", 1);
    run<2>("mad.wide.u32 (IMAD.WIDE)", 1);
    run<3>("mul.hi.u64", 1);
    run<4>("mad.lo.u64", 1);
    run<5>("add.cc+addc (2 IADD3)", 2);
    run<6>("fma.rn.f64 (DFMA)", 1);
    run<7>("wide + 2 add (3 instr)", 3);
    run<8>("mul.wide.u16", 1);
    run<9>("mul.lo + mul.hi u32", 2);
    cudaError_t e = cudaGetLastError();
    printf("status: %s\n", cudaGetErrorString(e));
    return 0;
}
