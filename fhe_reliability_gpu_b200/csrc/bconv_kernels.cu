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

// ---------------------------------------------------------------- tensor-core contraction (m <= 16)
// The contraction is a GEMM  out[i][k] = sum_j y[j][i] * hat[j][k]  with 60-bit integer entries.  Split both operands
// into BYTES: with k-index (j, a) = byte a of y_j and a Toeplitz-expanded constant matrix
//     B[(j, a)][(k, d)] = byte (d - a) of hat_jk   (0 outside 0 <= d - a < 8),
// the u8 x u8 -> s32 tensor-core product gives, per output, the 15 digit sums  S_d = sum_j sum_a y_j[a] hat_jk[d-a]
// (each < m * 8 * 255^2 < 2^23 for m <= 16) and  out = (sum_d S_d 2^(8d)) mod q_k : one 128-bit Barrett per output.
// y needs no repacking at all: the little-endian u64 words of y ARE the byte matrix (an A-fragment register is the low or
// high half of one y word).  mma.sync.m16n8k32 issues every 8 cycles per SM sub-partition on sm_100a (measured,
// profiles/microbench/imma_rates.cu): 16 of them per output and thread against 64 IMAD.WIDE at 4 cycles each before.
//
// Warp tile = 32 coefficients (two m-tiles) x 4 targets (eight n-tiles: columns 2t, 2t+1 of n-tile u hold digits
// 2u, 2u+1 of target base + t, so after the k loop lane (g, t) owns ALL 16 digits of target base + t for rows g and g + 8).
// The constant table is stored in fragment order [n-tile u][k-step s][target][8 lanes' worth] so that four consecutive
// targets are one contiguous, conflict-free 256-byte shared-memory read, for any first target.
constexpr int kTcWarps = 8;
#ifndef B200_TC_ABL
#define B200_TC_ABL 0
#endif

__device__ __forceinline__ void mma_u8(int (&c)[4], const u32 (&a)[4], const uint2 b) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y));
}

// sum_d S_d 2^(8d) as a 128-bit value; S_d < 2^23
__device__ __forceinline__ void digits_value(const u32 (&S)[16], u64& lo, u64& hi) {
    u64 W[4];
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const u32 x = S[4 * r] + (S[4 * r + 1] << 8), yv = S[4 * r + 2] + (S[4 * r + 3] << 8);   // each < 2^32
        W[r] = (u64)x + ((u64)yv << 16);
    }
    lo = W[0] + (W[1] << 32);
    hi = (W[1] >> 32) + W[2] + (W[3] << 32) + (lo < W[0]);
}

template <int KS>
__global__ void __launch_bounds__(kTcWarps * 32, 2)
    k_bconv_contract_tc(const u64* __restrict__ y, u64* __restrict__ out, const B200Mod* mods, const u32* dst_idx,
                        const uint2* __restrict__ btab /*[8][KS][K][8]*/, u32 N, u32 m, u32 K, u32 k0, u32 kcount, u32 per_cta) {
    extern __shared__ __align__(16) uint2 sb[];                 // [8][KS][kc][8] fragments, then [kc] (q, ratio0, ratio1, -)
    const u32 ka = k0 + blockIdx.y * per_cta;                   // first target of this CTA
    const u32 kend = (k0 + kcount < ka + per_cta) ? k0 + kcount : ka + per_cta;
    const u32 kc = kend - ka;
    {   // stage the constant fragments asynchronously: 8*KS rows of kc*64 contiguous bytes each
        const u32 row16 = kc * 4, total16 = 8 * KS * row16;
        const u32 sbase = (u32)__cvta_generic_to_shared(sb);
        for (u32 idx = threadIdx.x; idx < total16; idx += blockDim.x) {
            const u32 us = idx / row16, e = idx - us * row16;
            const char* src = reinterpret_cast<const char*>(btab + ((size_t)us * K + ka) * 8) + (size_t)e * 16;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sbase + idx * 16), "l"(src) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    ulonglong4* sq = reinterpret_cast<ulonglong4*>(sb + (size_t)8 * KS * kc * 8);
    for (u32 k = threadIdx.x; k < kc; k += blockDim.x) {
        const B200Mod& mk = mods[dst_idx[ka + k]];
        sq[k] = make_ulonglong4(mk.q, mk.ratio0, mk.ratio1, 0);
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    const u32* y32 = reinterpret_cast<const u32*>(y);
    const u32 n_wt = (N + 31) / 32;
    for (u32 wt = blockIdx.x * kTcWarps + warp; wt < n_wt; wt += gridDim.x * kTcWarps) {
        const u32 i0 = wt * 32;
        u32 a[2][KS][4];
#pragma unroll
        for (int mt = 0; mt < 2; mt++) {
            const u32 r0 = i0 + 16 * mt + g, r1 = r0 + 8;
#pragma unroll
            for (int s = 0; s < KS; s++) {
                const u32 j0 = 4 * s + (t >> 1), j1 = j0 + 2;
                a[mt][s][0] = (j0 < m && r0 < N) ? __ldg(y32 + ((size_t)j0 * N + r0) * 2 + (t & 1)) : 0;
                a[mt][s][1] = (j0 < m && r1 < N) ? __ldg(y32 + ((size_t)j0 * N + r1) * 2 + (t & 1)) : 0;
                a[mt][s][2] = (j1 < m && r0 < N) ? __ldg(y32 + ((size_t)j1 * N + r0) * 2 + (t & 1)) : 0;
                a[mt][s][3] = (j1 < m && r1 < N) ? __ldg(y32 + ((size_t)j1 * N + r1) * 2 + (t & 1)) : 0;
            }
        }
        for (u32 kb = 0; kb < kc; kb += 4) {
            int acc[2][8][4];
#pragma unroll
            for (int mt = 0; mt < 2; mt++)
#pragma unroll
                for (int u = 0; u < 8; u++) acc[mt][u][0] = acc[mt][u][1] = acc[mt][u][2] = acc[mt][u][3] = 0;
            u32 tb = kb + (g >> 1);                              // the target whose B columns this lane feeds
            if (tb > kc - 1) tb = kc - 1;
            const uint2* bp = sb + (size_t)tb * 8 + (g & 1) * 4 + t;
            // k-step outermost: the 16 accumulators are independent, so a dependent IMMA is 16 issues (not 2) behind its producer
#pragma unroll
            for (int s = 0; s < KS; s++) {
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    const uint2 b = bp[(size_t)(u * KS + s) * kc * 8];
#if B200_TC_ABL == 2          // timing ablation: no tensor work
                    acc[0][u][s & 3] += a[0][s][u & 3] ^ b.x;
                    acc[1][u][s & 3] += a[1][s][u & 3] ^ b.y;
#else
                    mma_u8(acc[0][u], a[0][s], b);
                    mma_u8(acc[1][u], a[1][s], b);
#endif
                }
            }
            // branch-free epilogue: the four outputs of this lane are independent chains the scheduler can interleave
            u32 tk = kb + t;                                     // the target whose digits this lane owns
            const bool live = tk < kc;
            if (!live) tk = kc - 1;
            const ulonglong4 mk = sq[tk];
            u64* orow = out + (size_t)(ka + tk - k0) * N;
            u64 res[2][2];
#pragma unroll
            for (int mt = 0; mt < 2; mt++) {
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    u32 S[16];
#pragma unroll
                    for (int d = 0; d < 16; d++) S[d] = (u32)acc[mt][d >> 1][(d & 1) + 2 * h];
                    u64 lo, hi;
#if B200_TC_ABL == 1          // timing ablation: no digit recombination / reduction
                    lo = 0;
#pragma unroll
                    for (int d = 0; d < 16; d++) lo ^= (u64)S[d] << (d & 7);
                    res[mt][h] = lo ^ mk.x;
#else
                    digits_value(S, lo, hi);
                    res[mt][h] = barrett128_raw(lo, hi, mk.x, mk.y, mk.z);
#endif
                }
            }
#pragma unroll
            for (int mt = 0; mt < 2; mt++) {
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const u32 row = i0 + 16 * mt + 8 * h + g;
                    if (live && row < N) orow[row] = res[mt][h];
                }
            }
        }
    }
}

template <int KS>
static void launch_tc(const u64* y, u64* out, const B200Mod* mods, const u32* dst_idx, const uint2* btab, u32 N, u32 m, u32 K, u32 k0,
                      u32 kcount, cudaStream_t s) {
    const size_t per_target = (size_t)KS * 8 * 8 * sizeof(uint2) + sizeof(ulonglong4);   // fragment + constant bytes per target
    const size_t budget = 100 * 1024;                                      // two CTAs per SM
    u32 ny = (u32)((kcount * per_target + budget - 1) / budget);
    u32 per_cta = ((kcount + ny - 1) / ny + 3) & ~3u;
    ny = (kcount + per_cta - 1) / per_cta;
    const u32 blocks = ((N + 31) / 32 + kTcWarps - 1) / kTcWarps;
    u32 gx = 296 / ny;
    if (gx < 1) gx = 1;
    if (gx > blocks) gx = blocks;
    const u32 rounds = (blocks + gx - 1) / gx;
    gx = (blocks + rounds - 1) / rounds;
    const size_t smem = per_target * (per_cta < kcount ? per_cta : kcount);
    cudaFuncSetAttribute(k_bconv_contract_tc<KS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget + 4096);
    k_bconv_contract_tc<KS><<<dim3(gx, ny), kTcWarps * 32, smem, s>>>(y, out, mods, dst_idx, btab, N, m, K, k0, kcount, per_cta);
}

// btab != nullptr and m <= 16: tensor-core path
int launch_bconv_contract_tc(const u64* y, u64* out, const B200Mod* mods, const u32* dst_idx, const uint2* btab, u32 N, u32 m, u32 K,
                             u32 k0, u32 kcount, cudaStream_t s) {
    if (!kcount) return 0;
    switch ((m + 3) / 4) {
        case 1: launch_tc<1>(y, out, mods, dst_idx, btab, N, m, K, k0, kcount, s); break;
        case 2: launch_tc<2>(y, out, mods, dst_idx, btab, N, m, K, k0, kcount, s); break;
        case 3: launch_tc<3>(y, out, mods, dst_idx, btab, N, m, K, k0, kcount, s); break;
        case 4: launch_tc<4>(y, out, mods, dst_idx, btab, N, m, K, k0, kcount, s); break;
        default: return -1;
    }
    return 0;
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
