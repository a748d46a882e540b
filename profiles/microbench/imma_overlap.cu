// Does the legacy IMMA (mma.sync u8) path on sm_100a overlap with ALU/FMA-pipe instructions of OTHER warps / the same
// warp?  Per loop iteration: 8 independent IMMAs + X independent integer ops (8 separate chains), 4 warps per SMSP.
// overlapped: cycles/iter ~ max(64, X * c);  additive: ~ 64 + X * c.
#include <cstdio>
#include <cuda_runtime.h>

template <int X, int KIND, int NIMMA>
__global__ void k(int iters, int* out, unsigned long long* cyc) {
    int acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; i++) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0;
    unsigned a0 = threadIdx.x * 0x01010101u, a1 = a0 ^ 0x5a5a5a5au, a2 = a0 + 7, a3 = a1 + 9, b0 = a0 * 3, b1 = a1 * 5;
    unsigned x[8];
    unsigned long long w[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { x[i] = threadIdx.x + i; w[i] = x[i]; }
    __syncthreads();
    unsigned long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (i < NIMMA)
                asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+r"(acc[i][0]), "+r"(acc[i][1]), "+r"(acc[i][2]), "+r"(acc[i][3])
                             : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
#pragma unroll
            for (int j = 0; j < X / 8; j++) {
                if (KIND == 0) asm volatile("add.u32 %0, %0, %1;" : "+r"(x[(i + j) & 7]) : "r"(a1));
                else asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[(i + j) & 7]) : "r"(a0), "r"(a1));
            }
        }
    }
    unsigned long long t1 = clock64();
    int s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += acc[i][0] + acc[i][1] + acc[i][2] + acc[i][3] + x[i] + (int)w[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int X, int KIND, int NIMMA>
void run(const char* name) {
    int* out; unsigned long long* cyc;
    const int threads = 512, blocks = 148, iters = 2048;   // 4 warps per SMSP
    cudaMalloc(&out, sizeof(int) * threads * blocks); cudaMalloc(&cyc, 8);
    k<X, KIND, NIMMA><<<blocks, threads>>>(16, out, cyc);
    k<X, KIND, NIMMA><<<blocks, threads>>>(iters, out, cyc);
    cudaDeviceSynchronize();
    unsigned long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-10s imma/iter %d  other/iter %3d : %.1f cycles per iteration per warp, %.1f per SMSP-iteration (4 warps)  %s\n", name, NIMMA, X,
           (double)c / iters, (double)c / iters / 4, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<0, 0, 8>("imma");
    run<16, 0, 8>("imma+IADD"); run<32, 0, 8>("imma+IADD"); run<64, 0, 8>("imma+IADD"); run<128, 0, 8>("imma+IADD");
    run<64, 0, 0>("IADD only"); run<128, 0, 0>("IADD only");
    run<16, 1, 8>("imma+WIDE"); run<32, 1, 8>("imma+WIDE"); run<64, 1, 8>("imma+WIDE");
    run<32, 1, 0>("WIDE only"); run<64, 1, 0>("WIDE only");
    return 0;
}
