// Fast RNS base conversion (ModUp), HPS / SEAL BaseConverter::fast_convert_array
// semantics (spec rfhe_framewk/src/baseConv.py:10-40, the GPU
// counterpart lives inside Phantom's relinearize/rotate, call sites
// reliability_test/dotprod_test.cu:114,146):
//     y_j   = x_j * (P/p_j)^-1 mod p_j                      (k_bconv_scale)
//     out_k = sum_j y_j * ((P/p_j) mod q_k)  mod q_k        (k_bconv_contract)
// The two steps are separate kernels because the limb-sharded ModUp (config C5)
// all-gathers y between them.  The contraction is integer-compute-bound
// (m*K multiply-accumulates per coefficient): every product is carry-free --
// y_j and the constants are split at 30 bits, the four partial products of up to
// 16 terms accumulate in 64-bit sums (4 IMAD.WIDE per term, no adds) -- and each
// output takes ONE 128-bit Barrett reduction.
#include "ctx.hpp"

namespace b200 {

// y[j][i] = x[j][i] * inv_j mod p_j ; one thread = two coefficients of one limb
__global__ void __launch_bounds__(256) k_bconv_scale(const u64* __restrict__ x, u64* __restrict__ y, const B200Mod* mods,
                                                     const u32* src_idx, const ulonglong2* inv, u32 N, u32 j0) {
    const u32 j = j0 + blockIdx.y;
    const B200Mod m = mods[src_idx[j]];
    const ulonglong2 w = inv[j];
    const size_t i2 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i2 * 2 >= N) return;
    const size_t off = ((size_t)blockIdx.y * N) / 2 + i2;
    const ulonglong2 v = reinterpret_cast<const ulonglong2*>(x)[off];
    const u64 nq = 0 - m.q;
    ulonglong2 r;
    r.x = canon2(shoup_mul(v.x, w.x, w.y, nq), m.q);
    r.y = canon2(shoup_mul(v.y, w.x, w.y, nq), m.q);
    reinterpret_cast<ulonglong2*>(y)[off] = r;
}

constexpr int kBcT = 4;          // outputs per accumulation tile
constexpr int kBcChunk = 16;     // terms per carry-free accumulation chunk ((2^30)^2 * 16 = 2^64)

struct Acc4 {
    u64 p00, p01, p10, p11;
};
__device__ __forceinline__ void acc4_value(const Acc4& a, u64& s0, u64& s1) {
    u64 t;
    s0 = a.p00; s1 = 0;
    t = a.p01 << 30; s0 += t; s1 += (a.p01 >> 34) + (s0 < t);
    t = a.p10 << 30; s0 += t; s1 += (a.p10 >> 34) + (s0 < t);
    t = a.p11 << 60; s0 += t; s1 += (a.p11 >> 4) + (s0 < t);
}

// out[k][i] = sum_j y[j][i] * hat[j][k] mod q_k for k in [k0, k0 + gridDim.y * kBcT).
// One thread = one coefficient and kBcT outputs; the y column of the CTA's 256
// coefficients is staged once in shared memory (split in 30-bit halves) and the
// constants are broadcast shared-memory reads.
__global__ void __launch_bounds__(256) k_bconv_contract(const u64* __restrict__ y, u64* __restrict__ out, const B200Mod* mods,
                                                        const u32* dst_idx, const u64* hat /*[K][m] packed halves*/, u32 N, u32 m,
                                                        u32 K, u32 k0, u32 kcount) {
    extern __shared__ u64 sm[];
    u64* ys = sm;                              // [m][256] packed halves (lo30 | hi30 << 32)
    u64* hs = sm + (size_t)m * 256;            // [kBcT][m]
    const u32 i = blockIdx.x * 256 + threadIdx.x;
    for (u32 j = 0; j < m; j++) {
        const u64 v = i < N ? y[(size_t)j * N + i] : 0;
        ys[j * 256 + threadIdx.x] = (v & 0x3fffffffull) | ((v >> 30) << 32);
    }
    const u32 n_tiles = (kcount + kBcT - 1) / kBcT;
    for (u32 tile = blockIdx.y; tile < n_tiles; tile += gridDim.y) {
        const u32 kt = k0 + tile * kBcT;           // first output of this tile
        __syncthreads();                           // ys staged / previous tile's constants no longer read
        for (u32 t = threadIdx.x; t < kBcT * m; t += 256) {
            const u32 k = kt + t / m;
            hs[t] = k < k0 + kcount ? hat[(size_t)k * m + t % m] : 0;
        }
        __syncthreads();
        u64 res[kBcT];
#pragma unroll
        for (int o = 0; o < kBcT; o++) res[o] = 0;
        for (u32 jc = 0; jc < m; jc += kBcChunk) {
            Acc4 acc[kBcT];
#pragma unroll
            for (int o = 0; o < kBcT; o++) acc[o] = Acc4{0, 0, 0, 0};
            const u32 je = jc + kBcChunk < m ? jc + kBcChunk : m;
            for (u32 j = jc; j < je; j++) {
                u32 y0, y1;
                unpack64(ys[j * 256 + threadIdx.x], y0, y1);
#pragma unroll
                for (int o = 0; o < kBcT; o++) {
                    u32 h0, h1;
                    unpack64(hs[o * m + j], h0, h1);
                    asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[o].p00) : "r"(y0), "r"(h0));
                    asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[o].p01) : "r"(y0), "r"(h1));
                    asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[o].p10) : "r"(y1), "r"(h0));
                    asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[o].p11) : "r"(y1), "r"(h1));
                }
            }
#pragma unroll
            for (int o = 0; o < kBcT; o++) {
                if (kt + o < k0 + kcount) {
                    const B200Mod mk = mods[dst_idx[kt + o]];
                    u64 s0, s1;
                    acc4_value(acc[o], s0, s1);
                    s0 += res[o];                  // previous chunks, already < q
                    s1 += (s0 < res[o]);
                    res[o] = barrett128(s0, s1, mk);
                }
            }
        }
        if (i < N) {
#pragma unroll
            for (int o = 0; o < kBcT; o++)
                if (kt + o < k0 + kcount) out[(size_t)(kt + o - k0) * N + i] = res[o];
        }
    }
}

int launch_bconv_scale(const u64* x, u64* y, const B200Mod* mods, const u32* src_idx, const ulonglong2* inv, u32 N, u32 j0, u32 count,
                       cudaStream_t s) {
    if (!count) return 0;
    dim3 grid((N / 2 + 255) / 256, count);
    k_bconv_scale<<<grid, 256, 0, s>>>(x, y, mods, src_idx, inv, N, j0);
    return 0;
}

int launch_bconv_contract(const u64* y, u64* out, const B200Mod* mods, const u32* dst_idx, const u64* hat, u32 N, u32 m, u32 K, u32 k0,
                          u32 kcount, cudaStream_t s) {
    if (!kcount) return 0;
    const size_t smem = sizeof(u64) * ((size_t)m * 256 + (size_t)kBcT * m);
    if (smem > 200 * 1024) return -1;
    cudaFuncSetAttribute(k_bconv_contract, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    // y is staged once per CTA; split the output tiles over grid.y only when there are too few CTAs to fill 148 SMs
    const u32 nx = (N + 255) / 256, n_tiles = (kcount + kBcT - 1) / kBcT;
    u32 ny = nx >= 592 ? 1 : (592 + nx - 1) / nx;
    if (ny > n_tiles) ny = n_tiles;
    dim3 grid(nx, ny);
    k_bconv_contract<<<grid, 256, smem, s>>>(y, out, mods, dst_idx, hat, N, m, K, k0, kcount);
    return 0;
}

}  // namespace b200
