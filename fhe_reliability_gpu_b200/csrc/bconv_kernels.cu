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

// ---- checksum-carrying conversion (spec rfhe_framewk/src/baseConv.py:42-98 bConv_ecc) ----------------
// The prototype appends e_j = sum_i x_j[i] mod p_j to every source limb, converts it with the data and compares, per
// term, the converted checksum with the sum of the converted data.  Here the same linearity is checked per KERNEL with
// sums accumulated where the data already is in registers (no extra pass over memory):
//   scale:     sum_i y_j[i] == inv_j * sum_i x_j[i]  (mod p_j)                      per source limb j
//   contract:  sum_i out_k[i] == sum_j hat_jk * (sum_i y_j[i])  (mod q_k)           per target limb k
//   in between: the y sums at the producer's store equal the y sums at the consumer's load (memory / transfer fault)
// Sums are exact integers kept as separate sums of the low and high 32-bit halves (each < 2^32 N); reduction mod the
// moduli happens once, in k_bconv_prot_finalize.
// BconvInj / BconvInjList (device copy of b200fhe_bconv_inject): ctx.hpp, shared with bconv_tc5.cu

__device__ __forceinline__ u64 warp_sum64(u64 v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ void red_add64(u64* p, u64 v) { asm volatile("red.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }

// counters c1 [m][4] = {sum x lo32, sum x hi32, sum y lo32, sum y hi32}
__global__ void __launch_bounds__(256) k_bconv_scale_prot(const u64* __restrict__ x, u64* __restrict__ y, const B200Mod* mods,
                                                          const u32* src_idx, const ulonglong2* inv, u32 N, u32 j0, u64* c1,
                                                          const __grid_constant__ BconvInjList inj) {
    const u32 j = j0 + blockIdx.y;
    const B200Mod m = mods[src_idx[j]];
    const ulonglong2 w = inv[j];
    const u64 nq = 0 - m.q;
    u64 sxl = 0, sxh = 0, syl = 0, syh = 0;
    constexpr int kPer = 4;                                   // coefficient pairs per thread: amortises the reductions
    ulonglong2 v[kPer];
    size_t idx[kPer];
#pragma unroll
    for (int r = 0; r < kPer; r++) {
        idx[r] = ((size_t)blockIdx.x * kPer + r) * blockDim.x + threadIdx.x;
        v[r] = idx[r] * 2 < N ? reinterpret_cast<const ulonglong2*>(x)[((size_t)blockIdx.y * N) / 2 + idx[r]] : make_ulonglong2(0, 0);
    }
#pragma unroll
    for (int r = 0; r < kPer; r++) {
        if (idx[r] * 2 >= N) continue;
        ulonglong2 o;
        o.x = canon2(shoup_mul(v[r].x, w.x, w.y, nq), m.q);
        o.y = canon2(shoup_mul(v[r].y, w.x, w.y, nq), m.q);
        for (u32 e = 0; e < inj.n; e++)                       // a faulty multiplier: the corrupted product is what gets stored AND summed
            if (inj.e[e].stage == 1 && inj.e[e].limb == j) {
                if (inj.e[e].coeff == idx[r] * 2) o.x ^= inj.e[e].mask;
                if (inj.e[e].coeff == idx[r] * 2 + 1) o.y ^= inj.e[e].mask;
            }
        reinterpret_cast<ulonglong2*>(y)[((size_t)blockIdx.y * N) / 2 + idx[r]] = o;
        sxl += (v[r].x & 0xffffffffull) + (v[r].y & 0xffffffffull); sxh += (v[r].x >> 32) + (v[r].y >> 32);
        syl += (o.x & 0xffffffffull) + (o.y & 0xffffffffull); syh += (o.x >> 32) + (o.y >> 32);
    }
    // one reduction per CTA and counter (per-warp reductions to the same four addresses serialise in L2: 71 us instead of 8)
    __shared__ u64 part[8][4];
    sxl = warp_sum64(sxl); sxh = warp_sum64(sxh); syl = warp_sum64(syl); syh = warp_sum64(syh);
    if ((threadIdx.x & 31) == 0) {
        u64* pw = part[threadIdx.x >> 5];
        pw[0] = sxl; pw[1] = sxh; pw[2] = syl; pw[3] = syh;
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        u64 v = 0;
#pragma unroll
        for (int w8 = 0; w8 < 8; w8++) v += part[w8][threadIdx.x];
        red_add64(c1 + (size_t)j * 4 + threadIdx.x, v);
    }
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

// PROT: also accumulates the checksum sums: cin [m][2] = sums of the low / high halves of the y words as loaded (grid.y
// slice 0 only), cout [K][2] = sums of the low / high halves of the outputs as stored (per-CTA shared-memory
// accumulators, one global reduction per CTA and target).
template <int KS, bool PROT>
__global__ void __launch_bounds__(kTcWarps * 32, 2)
    k_bconv_contract_tc(const u64* __restrict__ y, u64* __restrict__ out, const B200Mod* mods, const u32* dst_idx,
                        const uint2* __restrict__ btab /*[8][KS][K][8]*/, u32 N, u32 m, u32 K, u32 k0, u32 kcount, u32 per_cta,
                        u64* cin, u64* cout, const __grid_constant__ BconvInjList inj) {
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
    // PROT: per-warp checksum accumulators (plain read-add-write by one lane per entry: 64-bit shared atomics are CAS loops)
    u64* so = reinterpret_cast<u64*>(sq + kc);                   // [kTcWarps][kc][2] output sums
    u64* si = so + (size_t)kTcWarps * kc * 2;                    // [kTcWarps][16][2] input sums
    for (u32 k = threadIdx.x; k < kc; k += blockDim.x) {
        const B200Mod& mk = mods[dst_idx[ka + k]];
        sq[k] = make_ulonglong4(mk.q, mk.ratio0, mk.ratio1, 0);
    }
    if (PROT)
        for (u32 k = threadIdx.x; k < kTcWarps * (kc * 2 + 32); k += blockDim.x) so[k] = 0;
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
        if (PROT && blockIdx.y == 0 && cin) {
            // every (source limb, row, half) is held by exactly one fragment register of the warp: a[.][s][r] is the
            // (t & 1) half of y[4s + (t >> 1) + 2 (r >> 1)][row]; rows >= N and limbs >= m were loaded as 0
#pragma unroll
            for (int s2 = 0; s2 < KS; s2++)
#pragma unroll
                for (int jr = 0; jr < 2; jr++) {
                    u64 v = (u64)a[0][s2][2 * jr] + a[0][s2][2 * jr + 1] + a[1][s2][2 * jr] + a[1][s2][2 * jr + 1];
                    v += __shfl_xor_sync(0xffffffffu, v, 4);
                    v += __shfl_xor_sync(0xffffffffu, v, 8);
                    v += __shfl_xor_sync(0xffffffffu, v, 16);
                    const u32 jj = 4 * s2 + (t >> 1) + 2 * jr;
                    if (g == 0) si[(warp * 16 + jj) * 2 + (t & 1)] += v;
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
            if (PROT) {
                for (u32 e = 0; e < inj.n; e++)                   // a faulty reduction: the corrupted output is stored AND summed
                    if (inj.e[e].stage == 2 && live && inj.e[e].limb == ka + tk) {
#pragma unroll
                        for (int mt = 0; mt < 2; mt++)
#pragma unroll
                            for (int h = 0; h < 2; h++)
                                if (inj.e[e].coeff == i0 + 16 * mt + 8 * h + g) res[mt][h] ^= inj.e[e].mask;
                    }
                u64 sl = 0, sh = 0;
#pragma unroll
                for (int mt = 0; mt < 2; mt++)
#pragma unroll
                    for (int h = 0; h < 2; h++)
                        if (live && i0 + 16 * mt + 8 * h + g < N) { sl += res[mt][h] & 0xffffffffull; sh += res[mt][h] >> 32; }
                // sum over the 8 lanes that own the same target (same t).  (REDUX with the four per-t lane masks was measured
                // slower than these shuffles: 85 vs 75 us per launch.)
                sl += __shfl_xor_sync(0xffffffffu, sl, 4); sh += __shfl_xor_sync(0xffffffffu, sh, 4);
                sl += __shfl_xor_sync(0xffffffffu, sl, 8); sh += __shfl_xor_sync(0xffffffffu, sh, 8);
                sl += __shfl_xor_sync(0xffffffffu, sl, 16); sh += __shfl_xor_sync(0xffffffffu, sh, 16);
                if (g == 0 && live) {
                    u64* pw = so + ((size_t)warp * kc + tk) * 2;
                    pw[0] += sl;
                    pw[1] += sh;
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
    if (PROT) {
        __syncthreads();
        for (u32 k = threadIdx.x; k < 2 * kc + 32; k += blockDim.x) {
            u64 v = 0;
            if (k < 2 * kc) {
#pragma unroll
                for (int w8 = 0; w8 < kTcWarps; w8++) v += so[(size_t)w8 * kc * 2 + k];
                red_add64(cout + (size_t)ka * 2 + k, v);
            } else if (blockIdx.y == 0 && cin) {
                const u32 e = k - 2 * kc;
#pragma unroll
                for (int w8 = 0; w8 < kTcWarps; w8++) v += si[w8 * 32 + e];
                if (e / 2 < m) red_add64(cin + e, v);
            }
        }
    }
}

template <int KS, bool PROT>
static void launch_tc(const u64* y, u64* out, const B200Mod* mods, const u32* dst_idx, const uint2* btab, u32 N, u32 m, u32 K, u32 k0,
                      u32 kcount, cudaStream_t s, u64* cin, u64* cout, const BconvInjList& inj) {
    // fragment + constant (+ checksum accumulator) bytes per target
    const size_t per_target = (size_t)KS * 8 * 8 * sizeof(uint2) + sizeof(ulonglong4) + (PROT ? kTcWarps * 2 * sizeof(u64) : 0);
    const size_t fixed = PROT ? kTcWarps * 32 * sizeof(u64) : 0;
    const size_t budget = 108 * 1024;                                      // two CTAs per SM
    u32 ny = (u32)((kcount * per_target + budget - 1) / budget);
    u32 per_cta = ((kcount + ny - 1) / ny + 3) & ~3u;
    ny = (kcount + per_cta - 1) / per_cta;
    const u32 blocks = ((N + 31) / 32 + kTcWarps - 1) / kTcWarps;
    u32 gx = 296 / ny;
    if (gx < 1) gx = 1;
    if (gx > blocks) gx = blocks;
    const u32 rounds = (blocks + gx - 1) / gx;
    gx = (blocks + rounds - 1) / rounds;
    const size_t smem = per_target * (per_cta < kcount ? per_cta : kcount) + fixed;
    (void)set_smem_once<k_bconv_contract_tc<KS, PROT>>(budget + 4096);     // on failure the launch below fails and the caller's cudaGetLastError reports it
    k_bconv_contract_tc<KS, PROT><<<dim3(gx, ny), kTcWarps * 32, smem, s>>>(y, out, mods, dst_idx, btab, N, m, K, k0, kcount, per_cta,
                                                                              cin, cout, inj);
}

template <bool PROT>
static int launch_tc_ks(const u64* y, u64* out, const B200Mod* mods, const u32* dst_idx, const uint2* btab, u32 N, u32 m, u32 K, u32 k0,
                        u32 kcount, cudaStream_t s, u64* cin, u64* cout, const BconvInjList& inj) {
    if (!kcount) return 0;
    switch ((m + 3) / 4) {
        case 1: launch_tc<1, PROT>(y, out, mods, dst_idx, btab, N, m, K, k0, kcount, s, cin, cout, inj); break;
        case 2: launch_tc<2, PROT>(y, out, mods, dst_idx, btab, N, m, K, k0, kcount, s, cin, cout, inj); break;
        case 3: launch_tc<3, PROT>(y, out, mods, dst_idx, btab, N, m, K, k0, kcount, s, cin, cout, inj); break;
        case 4: launch_tc<4, PROT>(y, out, mods, dst_idx, btab, N, m, K, k0, kcount, s, cin, cout, inj); break;
        default: return -1;
    }
    return 0;
}

// btab != nullptr and m <= 16: tensor-core path
int launch_bconv_contract_tc(const u64* y, u64* out, const B200Mod* mods, const u32* dst_idx, const uint2* btab, u32 N, u32 m, u32 K,
                             u32 k0, u32 kcount, cudaStream_t s) {
    BconvInjList none;
    none.n = 0;
    return launch_tc_ks<false>(y, out, mods, dst_idx, btab, N, m, K, k0, kcount, s, nullptr, nullptr, none);
}

// One warp per checked limb.  counters: c1 [m][4] | cin [m][2] | cout [K][2]; two sets alternate between calls.
// verdict [m + K]: source limb j: bit 0 = scale check failed, bit 1 = y changed between the two kernels;
//                  target limb k (at m + k): bit 0 = contraction check failed.  sums (optional) [m + K][2] = (lhs, rhs).
__global__ void k_bconv_prot_finalize(u64* counters, u64* next_counters, const B200Mod* mods, const u32* src_idx, const u32* dst_idx, const ulonglong2* inv,
                                      const u64* hat /*[K][m] packed 30-bit halves*/, u32 m, u32 K, u32* verdict, u64* sums) {
    const u32 i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;     // one warp per check
    // the NEXT call's counter set is cleared here (the two sets alternate), which saves a clearing launch per call
    for (u32 z = blockIdx.x * blockDim.x + threadIdx.x; z < 6 * m + 2 * K; z += gridDim.x * blockDim.x) next_counters[z] = 0;
    if (i >= m + K) return;
    u64* c1 = counters;
    u64* cin = counters + (size_t)4 * m;
    u64* cout = cin + (size_t)2 * m;
    // (lo32 sum) + (hi32 sum) 2^32 as a 128-bit integer, reduced mod q
    auto value_mod = [](u64 sl, u64 sh, const B200Mod& mq) {
        const u64 lo = sl + (sh << 32);
        const u64 hi = (sh >> 32) + (lo < sl);
        return barrett128(lo, hi, mq);
    };
    u64 lhs, rhs;
    u32 v = 0;
    if (i < m) {
        const B200Mod mq = mods[src_idx[i]];
        lhs = value_mod(c1[4 * i + 2], c1[4 * i + 3], mq);                                    // sum y
        rhs = mulmod_barrett(value_mod(c1[4 * i + 0], c1[4 * i + 1], mq), inv[i].x, mq);      // inv * sum x
        if (lhs != rhs) v |= 1u;
        if (c1[4 * i + 2] != cin[2 * i] || c1[4 * i + 3] != cin[2 * i + 1]) v |= 2u;
    } else {
        const u32 k = i - m;
        const B200Mod mq = mods[dst_idx[k]];
        lhs = value_mod(cout[2 * k], cout[2 * k + 1], mq);                                    // sum out
        u64 term = 0;                                                                         // lane j: hat_jk * (sum y_j mod q_k)
        if (lane < m) {
            const u64 hp = hat[(size_t)k * m + lane];
            const u64 hv = (hp & 0x3fffffffull) + ((hp >> 32) << 30);
            term = mulmod_barrett(value_mod(cin[2 * lane], cin[2 * lane + 1], mq), hv, mq);
        }
        term = warp_sum64(term);                                                              // m <= 16 terms < 2^60 each
        rhs = barrett128(term, 0, mq);
        if (lhs != rhs) v |= 1u;
    }
    if (lane == 0) {
        verdict[i] = v;
        if (sums) { sums[2 * i] = lhs; sums[2 * i + 1] = rhs; }
    }
}
__global__ void k_bconv_xor(u64* y, size_t idx, u64 mask) { y[idx] ^= mask; }

// Protected ModUp of one digit on one GPU: scale -> [memory-fault injection] -> contraction -> finalize.
// inj: host array of n_inj entries {stage, limb, coeff, mask}; stage 1 = scale product, 2 = contraction output, 3 = y in memory.
// out_hi != nullptr: targets [0, k_split) go to out, targets [k_split, K) to out_hi (the key-switch writes the targets below and
// above a digit's own limbs into different rows); the second launch does not count the input sums again.
int launch_bconv_modup_protected(const u64* x, u64* y, u64* out, const B200Mod* mods, const u32* src_idx, const u32* dst_idx,
                                 const ulonglong2* inv, const u64* hat, const uint2* btab, u32 N, u32 m, u32 K, u64* counters,
                                 u64* next_counters, u32* verdict, u64* sums, const void* inj_host, u32 n_inj, cudaStream_t s, unsigned long long* launches,
                                 u64* out_hi, u32 k_split, const uint4* btab5) {
    if (!btab || m > 16 || n_inj > (u32)kBconvMaxInj) return -1;
    BconvInjList il;
    il.n = n_inj;
    for (u32 e = 0; e < n_inj; e++) il.e[e] = reinterpret_cast<const BconvInj*>(inj_host)[e];
    dim3 grid((N / 2 + 1023) / 1024, m);
    k_bconv_scale_prot<<<grid, 256, 0, s>>>(x, y, mods, src_idx, inv, N, 0, counters, il);
    unsigned long long nl = 1;
    for (u32 e = 0; e < n_inj; e++)
        if (il.e[e].stage == 3 && il.e[e].limb < m && il.e[e].coeff < N) {
            k_bconv_xor<<<1, 1, 0, s>>>(y, (size_t)il.e[e].limb * N + il.e[e].coeff, il.e[e].mask);
            nl++;
        }
    u64* cin = counters + (size_t)4 * m;
    u64* cout = counters + (size_t)6 * m;
    // tcgen05 contraction first (bconv_tc5.cu); the mma.sync kernel serves the shapes it declines
    auto contract = [&](u64* o, u32 k0, u32 kcount, u64* ci) {
        if (launch_bconv_contract_tc5_prot(y, o, mods, dst_idx, btab5, N, m, K, k0, kcount, s, ci, cout, il) == 0) return 0;
        return launch_tc_ks<true>(y, o, mods, dst_idx, btab, N, m, K, k0, kcount, s, ci, cout, il);
    };
    int rc;
    if (out_hi && k_split < K) {
        rc = k_split ? contract(out, 0, k_split, cin) : 0;
        if (!rc) rc = contract(out_hi, k_split, K - k_split, k_split ? nullptr : cin);
        nl += k_split ? 1 : 0;
    } else {
        rc = contract(out, 0, K, cin);
    }
    if (rc) return rc;
    k_bconv_prot_finalize<<<(m + K + 3) / 4, 128, 0, s>>>(counters, next_counters, mods, src_idx, dst_idx, inv, hat, m, K, verdict, sums);
    if (launches) *launches += nl + 2;
    return 0;
}

// ---- exact-CRT mode (motivation/baseConv.py:67-83 base_conv_fixed) --------------------------------------------------
// The fast conversion returns x + e P with e = floor(sum_j y_j / p_j) in [0, m).  The exact flavour takes e P out again.
// e has to be EXACT (a floating-point estimate of sum_j y_j / p_j cannot separate x = 0..few from e - tiny), so it is
// found in multi-word integers: S = sum_j y_j (P/p_j)  (m words: 60 m + log2 m <= 64 m bits) is compared with the
// multiples e P; then out_k -= (e P mod q_k).  One thread per coefficient; a mode, not the hot path (the key-switch
// wants the fast flavour, whose overshoot it absorbs as noise).
constexpr int kExactMaxM = 64;
__global__ void __launch_bounds__(128) k_bconv_exact_fix(const u64* __restrict__ y, u64* __restrict__ out, u32 N, u32 m, u32 K,
                                                         const u64* __restrict__ hatw /*[m][m]*/, const u64* __restrict__ ePw /*[m][m]*/,
                                                         const u64* __restrict__ ePq /*[K][m]*/, const B200Mod* mods, const u32* dst_idx) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    u64 S[kExactMaxM];
    for (u32 t = 0; t < m; t++) S[t] = 0;
    for (u32 j = 0; j < m; j++) {
        const u64 yj = y[(size_t)j * N + i];
        const u64* h = hatw + (size_t)j * m;
        u64 carry = 0;
        for (u32 t = 0; t < m; t++) {                 // S += yj * h, word by word
            const u64 lo = yj * h[t], hi = __umul64hi(yj, h[t]);
            u64 s = S[t] + lo;
            u64 c = s < lo;
            s += carry;
            c += s < carry;
            S[t] = s;
            carry = hi + c;                           // hi <= 2^64 - 2, c <= 2: cannot wrap past 2^64 given the bound on S
        }
    }
    u32 e = 0;
    for (u32 c = m - 1; c >= 1; c--) {                // largest e with e P <= S
        const u64* w = ePw + (size_t)c * m;
        bool ge = true;
        for (int t = (int)m - 1; t >= 0; t--) {
            if (S[t] != w[t]) { ge = S[t] > w[t]; break; }
        }
        if (ge) { e = c; break; }
    }
    if (e == 0) return;
    for (u32 k = 0; k < K; k++) {
        const u64 q = mods[dst_idx[k]].q, sub = ePq[(size_t)k * m + e];
        u64 v = out[(size_t)k * N + i];
        v = v >= sub ? v - sub : v + q - sub;
        out[(size_t)k * N + i] = v;
    }
}
int launch_bconv_exact_fix(const u64* y, u64* out, u32 N, u32 m, u32 K, const u64* hatw, const u64* ePw, const u64* ePq, const B200Mod* mods,
                           const u32* dst_idx, cudaStream_t s) {
    if (m > (u32)kExactMaxM) return -1;
    k_bconv_exact_fix<<<(N + 127) / 128, 128, 0, s>>>(y, out, N, m, K, hatw, ePw, ePq, mods, dst_idx);
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
    if (set_smem_once<k_bconv_contract>(200 * 1024) != cudaSuccess) return -1;
    // y is staged once per CTA; split the output tiles over grid.y only when there are too few CTAs to fill 148 SMs
    const u32 nx = (N + 255) / 256, n_tiles = (kcount + kBcT - 1) / kBcT;
    u32 ny = nx >= 592 ? 1 : (592 + nx - 1) / nx;
    if (ny > n_tiles) ny = n_tiles;
    dim3 grid(nx, ny);
    k_bconv_contract<<<grid, 256, smem, s>>>(y, out, mods, dst_idx, hat, N, m, K, k0, kcount);
    return 0;
}

}  // namespace b200
