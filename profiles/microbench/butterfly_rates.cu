// Register-only Harvey butterfly throughput on sm_100a for several decompositions
// of the 64-bit Shoup multiply.  Upper bound for the NTT kernels (no memory).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o butterfly_rates butterfly_rates.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
typedef unsigned int u32;

// V0: what the compiler makes of __umul64hi + two 64-bit low products
__device__ __forceinline__ u64 shoup_v0(u64 y, u64 w, u64 ws, u64 q, u64 nq) {
    (void)nq;
    return y * w - __umul64hi(y, ws) * q;
}
// V1: exact high word with 4 carry-free mad.wide, low words fused into one mad chain on -q
__device__ __forceinline__ u64 shoup_v1(u64 y, u64 w, u64 ws, u64 q, u64 nq) {
    (void)q;
    u32 yl = (u32)y, yh = (u32)(y >> 32), wl = (u32)w, wh = (u32)(w >> 32);
    u32 sl = (u32)ws, sh = (u32)(ws >> 32), nl = (u32)nq, nh = (u32)(nq >> 32);
    u64 p0, s1, s2, k, r;
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(p0) : "r"(yl), "r"(sl));
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(s1) : "r"(yl), "r"(sh), "l"(p0 >> 32));
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(s2) : "r"(yh), "r"(sl), "l"(s1 & 0xffffffffull));
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(k) : "r"(yh), "r"(sh), "l"(s1 >> 32));
    k += s2 >> 32;
    u32 kl = (u32)k, kh = (u32)(k >> 32);
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(r) : "r"(yl), "r"(wl));
    asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(r) : "r"(kl), "r"(nl));
    u32 hi = (u32)(r >> 32);
    hi += yl * wh + yh * wl + kl * nh + kh * nl;
    return (r & 0xffffffffull) | ((u64)hi << 32);
}
// V2: approximate high word (drops yl*sl): result in [0,3q)
__device__ __forceinline__ u64 shoup_v2(u64 y, u64 w, u64 ws, u64 q, u64 nq) {
    (void)q;
    u32 yl = (u32)y, yh = (u32)(y >> 32), wl = (u32)w, wh = (u32)(w >> 32);
    u32 sl = (u32)ws, sh = (u32)(ws >> 32), nl = (u32)nq, nh = (u32)(nq >> 32);
    u64 s1, s2, k, r;
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(s1) : "r"(yl), "r"(sh));
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(s2) : "r"(yh), "r"(sl), "l"(s1 & 0xffffffffull));
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(k) : "r"(yh), "r"(sh), "l"(s1 >> 32));
    k += s2 >> 32;
    u32 kl = (u32)k, kh = (u32)(k >> 32);
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(r) : "r"(yl), "r"(wl));
    asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(r) : "r"(kl), "r"(nl));
    u32 hi = (u32)(r >> 32);
    hi += yl * wh + yh * wl + kl * nh + kh * nl;
    return (r & 0xffffffffull) | ((u64)hi << 32);
}
// V3: like V2 but the quotient is a 32-bit estimate from the top words only when y < 2^32*q_hi ... (not valid in
// general; measures the floor of a 2-wide-product estimate for reference only)
__device__ __forceinline__ u64 shoup_v3(u64 y, u64 w, u64 ws, u64 q, u64 nq) {
    (void)q;
    u32 yl = (u32)y, yh = (u32)(y >> 32), wl = (u32)w, wh = (u32)(w >> 32);
    u32 sh = (u32)(ws >> 32), nl = (u32)nq, nh = (u32)(nq >> 32);
    u64 s1, k, r;
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(s1) : "r"(yl), "r"(sh));
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(k) : "r"(yh), "r"(sh), "l"(s1 >> 32));
    u32 kl = (u32)k, kh = (u32)(k >> 32);
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(r) : "r"(yl), "r"(wl));
    asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(r) : "r"(kl), "r"(nl));
    u32 hi = (u32)(r >> 32);
    hi += yl * wh + yh * wl + kl * nh + kh * nl;
    return (r & 0xffffffffull) | ((u64)hi << 32);
}

template <int V>
__device__ __forceinline__ u64 smul(u64 y, u64 w, u64 ws, u64 q, u64 nq) {
    if (V == 0) return shoup_v0(y, w, ws, q, nq);
    if (V == 1) return shoup_v1(y, w, ws, q, nq);
    if (V == 2) return shoup_v2(y, w, ws, q, nq);
    return shoup_v3(y, w, ws, q, nq);
}

#define ITERS 64
template <int V>
__global__ void __launch_bounds__(1024) k(u64* out, u64 q, u64 wseed, long long* cyc, int* smhist) {
    extern __shared__ u64 dummy_smem[];
    if (threadIdx.x == 0) { u32 sm; asm volatile("mov.u32 %0, %%smid;" : "=r"(sm)); atomicAdd(&smhist[sm], 1); }
    u64 v[16];
    const u64 nq = 0 - q, two_q = 2 * q;
    u64 w[4], ws[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        w[i] = (wseed * (threadIdx.x + 3 + i)) % q;
        ws[i] = (u64)((((unsigned __int128)w[i]) << 64) / q);
    }
#pragma unroll
    for (int e = 0; e < 16; e++) v[e] = (wseed * (e + 1) + threadIdx.x) % q;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int t = 0; t < 4; t++) {
            const int half = 8 >> t;
#pragma unroll
            for (int g = 0; g < (1 << t); g++) {
#pragma unroll
                for (int i = 0; i < half; i++) {
                    const int a = g * 2 * half + i, b = a + half;
                    const u64 T = smul<V>(v[b], w[t], ws[t], q, nq);
                    const u64 X = v[a];
                    v[a] = X + T;
                    v[b] = X - T + two_q;
                }
            }
        }
        // keep values bounded like a real phase start does (cheap, alu only)
#pragma unroll
        for (int e = 0; e < 16; e++) v[e] &= 0x0fffffffffffffffull;
    }
    long long t1 = clock64();
    u64 acc = 0;
#pragma unroll
    for (int e = 0; e < 16; e++) acc ^= v[e];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) atomicMax((unsigned long long*)cyc, (unsigned long long)(t1 - t0));
}

template <int V>
void run(const char* name) {
    u64* out;
    long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 8 * 2);
    cudaMallocManaged(&cyc, 8);
    const u64 q = 1152921504606584833ull;
    int* hist;
    cudaMallocManaged(&hist, 256 * 4);
    cudaFuncSetAttribute(k<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024);
    for (int threads = 128; threads <= 1024; threads += (threads < 256 ? 128 : 256)) {
        int mx = 0;
        for (int rep = 0; rep < 2; rep++) {
            *cyc = 0;
            for (int i = 0; i < 256; i++) hist[i] = 0;
            k<V><<<148, threads, 120 * 1024>>>(out, q, 0x9E3779B97F4A7C15ull, cyc, hist);
            cudaDeviceSynchronize();
            mx = 0;
            for (int i = 0; i < 256; i++) mx = hist[i] > mx ? hist[i] : mx;
        }
        if (mx != 1) printf("  (max CTAs on one SM = %d)\n", mx);
        double warp_bf_per_smsp = (double)ITERS * 32 * (threads / 32) / 4.0;
        printf("%-34s warps/SM=%2d  cycles=%8lld  cyc/warp-butterfly/SMSP=%.2f\n", name, threads / 32, *cyc,
               *cyc / warp_bf_per_smsp);
    }
    cudaFree(out);
    cudaFree(cyc);
}

int main() {
    run<0>("V0 compiler umul64hi");
    run<1>("V1 exact 6 wide + 4 imad");
    run<2>("V2 approx-hi 5 wide + 4 imad");
    run<3>("V3 (invalid) 4 wide + 4 imad");
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
