// Legacy tensor path on sm_100a: issue rate of mma.sync.m16n8k32 (u8 x u8 -> s32) per SM sub-partition, and how it
// overlaps with IMAD.WIDE / ALU work from the same warp.  Decides whether the dense base-conversion contraction
// (bconv_kernels.cu) is worth moving from 30-bit-split IMAD.WIDE products to byte-split tensor-core products.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o imma_rates imma_rates.cu && ./imma_rates
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int NACC, int MIX>
__global__ void k_imma(int iters, int* out, unsigned long long* cyc) {
    int acc[NACC][4];
#pragma unroll
    for (int i = 0; i < NACC; i++) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0;
    unsigned a0 = threadIdx.x * 0x01010101u, a1 = a0 ^ 0x5a5a5a5au, a2 = a0 + 7, a3 = a1 + 9, b0 = a0 * 3, b1 = a1 * 5;
    unsigned long long w = threadIdx.x;
    unsigned x = threadIdx.x | 1;
    __syncthreads();
    unsigned long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i++) {
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+r"(acc[i][0]), "+r"(acc[i][1]), "+r"(acc[i][2]), "+r"(acc[i][3])
                         : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            if (MIX == 1) {   // 2 IMAD.WIDE per mma
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w) : "r"(x), "r"(a0));
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w) : "r"(x), "r"(a1));
            }
            if (MIX == 2) {   // 4 ALU ops per mma
                asm volatile("shf.l.wrap.b32 %0, %0, %1, 3;" : "+r"(x) : "r"(a0));
                asm volatile("add.u32 %0, %0, %1;" : "+r"(x) : "r"(a1));
                asm volatile("xor.b32 %0, %0, %1;" : "+r"(x) : "r"(a2));
                asm volatile("add.u32 %0, %0, %1;" : "+r"(x) : "r"(a3));
            }
        }
    }
    unsigned long long t1 = clock64();
    int s = (int)w + (int)x;
#pragma unroll
    for (int i = 0; i < NACC; i++) s += acc[i][0] + acc[i][1] + acc[i][2] + acc[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int NACC, int MIX>
void run(const char* name, int warps_per_smsp) {
    int* out; unsigned long long* cyc;
    const int threads = 128 * warps_per_smsp, blocks = 148, iters = 4096;
    cudaMalloc(&out, sizeof(int) * threads * blocks);
    cudaMalloc(&cyc, 8);
    k_imma<NACC, MIX><<<blocks, threads>>>(16, out, cyc);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k_imma<NACC, MIX><<<blocks, threads>>>(iters, out, cyc);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    unsigned long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    const double mmas_per_smsp = (double)iters * NACC * warps_per_smsp;
    const double tops = 2.0 * 16 * 8 * 32 * iters * NACC * (threads / 32) * blocks / (ms * 1e-3) / 1e12;
    printf("%-28s warps/SMSP %d  acc %d : %.2f cycles per mma per SMSP, %.1f TOPS chip-wide (%s)\n", name, warps_per_smsp, NACC,
           (double)c / mmas_per_smsp, tops, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<8, 0>("imma only", 1);
    run<8, 0>("imma only", 2);
    run<8, 0>("imma only", 4);
    run<16, 0>("imma only", 2);
    run<8, 1>("imma + 2 IMAD.WIDE", 2);
    run<8, 1>("imma + 2 IMAD.WIDE", 4);
    run<8, 2>("imma + 4 ALU", 2);
    run<8, 2>("imma + 4 ALU", 4);
    return 0;
}
