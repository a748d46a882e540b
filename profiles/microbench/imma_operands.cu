// IMMA (mma.sync u8 m16n8k32) issue rate on sm_100a as a function of operand variety: the same A/B registers every
// time (operand-reuse caches hit) against the access pattern of k_bconv_contract_tc (16 accumulators, A alternating
// between two m-tiles and four k-steps, a different B fragment for every pair of IMMAs).
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void mma(int (&c)[4], const unsigned (&a)[4], unsigned b0, unsigned b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

template <int MODE>
__global__ void __launch_bounds__(512, 1) k(int iters, int* out, unsigned long long* cyc) {
    int acc[2][8][4];
#pragma unroll
    for (int m = 0; m < 2; m++)
#pragma unroll
        for (int u = 0; u < 8; u++) acc[m][u][0] = acc[m][u][1] = acc[m][u][2] = acc[m][u][3] = 0;
    unsigned a[2][4][4], b[8][2];
#pragma unroll
    for (int m = 0; m < 2; m++)
#pragma unroll
        for (int s = 0; s < 4; s++)
#pragma unroll
            for (int r = 0; r < 4; r++) a[m][s][r] = threadIdx.x * 0x01010101u + m * 77 + s * 13 + r;
#pragma unroll
    for (int u = 0; u < 8; u++) { b[u][0] = threadIdx.x * 3 + u; b[u][1] = threadIdx.x * 5 + u; }
    extern __shared__ uint2 sm[];
    if (MODE == 4) for (int i = threadIdx.x; i < 32 * 12 * 32; i += blockDim.x) sm[i] = make_uint2(i, i * 7);
    __syncthreads();
    unsigned long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int s = 0; s < 4; s++)
#pragma unroll
            for (int u = 0; u < 8; u++) {
                if (MODE == 0) { mma(acc[0][u], a[0][0], b[0][0], b[0][1]); mma(acc[1][u], a[0][0], b[0][0], b[0][1]); }   // same operands
                if (MODE == 1) { mma(acc[0][u], a[0][s], b[u][0], b[u][1]); mma(acc[1][u], a[1][s], b[u][0], b[u][1]); }   // kernel pattern
                if (MODE == 2) { mma(acc[0][u], a[0][s], b[u][0], b[u][1]); mma(acc[1][u], a[0][s], b[u][0], b[u][1]); }   // A fixed per k-step
                if (MODE == 4) {   // kernel pattern with the B fragment loaded from shared memory right before its two IMMAs
                    const uint2 bb = sm[((u * 4 + s) * 12 + (it & 7)) * 32 + (threadIdx.x & 31)];
                    mma(acc[0][u], a[0][s], bb.x, bb.y); mma(acc[1][u], a[1][s], bb.x, bb.y);
                }
                if (MODE == 3) { mma(acc[0][u], a[0][s], b[0][0], b[0][1]); mma(acc[1][u], a[1][s], b[0][0], b[0][1]); }   // B fixed
            }
    }
    unsigned long long t1 = clock64();
    int sres = 0;
#pragma unroll
    for (int m = 0; m < 2; m++)
#pragma unroll
        for (int u = 0; u < 8; u++) sres += acc[m][u][0] + acc[m][u][1] + acc[m][u][2] + acc[m][u][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = sres;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int MODE>
void run(const char* name, int warps_per_smsp) {
    int* out; unsigned long long* cyc;
    const int threads = 128 * warps_per_smsp, blocks = 148, iters = 1024;
    cudaMalloc(&out, sizeof(int) * threads * blocks); cudaMalloc(&cyc, 8);
    const size_t smem = MODE == 4 ? 32 * 12 * 32 * sizeof(uint2) : 0;
    cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k<MODE><<<blocks, threads, smem>>>(8, out, cyc);
    k<MODE><<<blocks, threads, smem>>>(iters, out, cyc);
    cudaDeviceSynchronize();
    unsigned long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-34s warps/SMSP %d : %.2f cycles per IMMA per SMSP  %s\n", name, warps_per_smsp, (double)c / iters / 64 / warps_per_smsp,
           cudaGetErrorString(cudaGetLastError()));
    cudaFree(out); cudaFree(cyc);
}

int main() {
    for (int w = 1; w <= 4; w *= 2) {
        run<0>("same A and B every time", w);
        run<1>("kernel pattern (A alt., B per pair)", w);
        run<2>("A per k-step, B per pair", w);
        run<3>("A alternating, B fixed", w);
        run<4>("kernel pattern, B from shared", w);
    }
    return 0;
}
