// Fused ABFT weighted checksum of the forward NTT (config C4).
//
// Identity (rfhe_framewk/src/negaclic_ntt.py:130-149 ecc_check):
//     sum_i w_i a_i  ==  sum_k omega_k NTT(a)_k   (mod q),   omega = N^-1 NTT_{psi^-1}(w)
// with the prototype's small weights w = generate_weights(N) (:8-13).  The left
// side is accumulated while the FIRST pass loads the coefficients, the right
// side while the LAST pass stores the results -- no extra pass over the data
// (north_star: "checksum accumulation fused into the final stage").  A fault in
// any butterfly in between breaks the identity unless its effect on the sum is a
// multiple of q (miss probability ~ 1/q, as in relia_ntt_sim's batch checks).
//
// Cost control: every weighted product is carry-free.  w_i < 2^12, so
// w_i * lo32(a) and w_i * hi32(a) are accumulated by two IMAD.WIDE into two
// 64-bit sums; omega_k and the canonical outputs are split at 30 bits so the four
// partial products of 16 coefficients fit 64-bit sums.  One Barrett reduction
// partial products of 16 coefficients fit 64-bit sums.  The per-thread sums are committed
// unreduced (26-bit digits, one REDUX per digit per warp, the warps' digit sums added in shared memory,
// one RED per digit per CTA); the single modular reduction per limb happens in k_abft_finalize.
#pragma once
#include "modarith.cuh"

struct AbftIn {
    u64 lo = 0, hi = 0;
    __device__ __forceinline__ void add(u32 w, u64 x) {
        u32 xl, xh;
        unpack64(x, xl, xh);
        asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(lo) : "r"(w), "r"(xl));
        asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(hi) : "r"(w), "r"(xh));
    }
    // lo + hi 2^32 as a 128-bit number (< 2^92 for up to 32 terms of w < 2^12)
    __device__ __forceinline__ void value(u64& s0, u64& s1) const {
        s0 = lo + (hi << 32);
        s1 = (hi >> 32) + (s0 < lo);
    }
};

struct AbftOut {
    u64 p00 = 0, p01 = 0, p10 = 0, p11 = 0;
    // om = omega0 | omega1 << 32 (30-bit halves), x canonical (< 2^60)
    __device__ __forceinline__ void add(u64 om, u64 x) {
        u32 o0, o1;
        unpack64(om, o0, o1);
        const u32 x0 = (u32)x & 0x3fffffffu, x1 = (u32)(x >> 30);
#ifdef B200_ABL_ABFT_NOOUTMUL        // timing ablation: weights still loaded and consumed, no multiplier-pipe work
        p00 += (u64)(o0 ^ x0); p01 += (u64)(o1 ^ x1);
        return;
#endif
        asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(p00) : "r"(o0), "r"(x0));
        asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(p01) : "r"(o0), "r"(x1));
        asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(p10) : "r"(o1), "r"(x0));
        asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(p11) : "r"(o1), "r"(x1));
    }
    // p00 + (p01 + p10) 2^30 + p11 2^60  < 2^125
    __device__ __forceinline__ void value(u64& s0, u64& s1) const {
        u64 t;
        s0 = p00; s1 = 0;
        t = p01 << 30; s0 += t; s1 += (p01 >> 34) + (s0 < t);
        t = p10 << 30; s0 += t; s1 += (p10 >> 34) + (s0 < t);
        t = p11 << 60; s0 += t; s1 += (p11 >> 4) + (s0 < t);
    }
};

// ---- the same two accumulators on the FP64 pipe ----------------------------------------------------------------------
// The integer-multiply pipe is the binding unit of the transform; the FP64 pipe (one DFMA per 2 cycles per
// sub-partition, its own pipe) is idle.  Integer-valued doubles below 2^53 multiply and add exactly, so the weighted sums
// can be carried there at ZERO multiplier-pipe cost: a 32-bit piece p becomes the double 2^52 + p by pairing it with the
// high word 0x43300000 (no instruction), one DADD removes the 2^52, DFMA accumulates.
//   input side : w (< 2^12) x the two 32-bit halves of the word: 44-bit products, up to 32 terms      (< 2^49)
//   output side: omega's 30-bit halves x the three 20-bit pieces of the canonical output: 50-bit products, 8 terms per
//                accumulator (two accumulator sets, even / odd outputs)                                (< 2^53)
// The sums are converted back to integers once per thread; from there on the path is the integer one (same 128-bit value).
__device__ __forceinline__ double abft_dbl(u32 piece) { return __hiloint2double(0x43300000, (int)piece) - 4503599627370496.0; }

struct AbftInF {
    double lo = 0.0, hi = 0.0;
    __device__ __forceinline__ void add(double w, u64 x) {
        u32 xl, xh;
        unpack64(x, xl, xh);
        lo = fma(w, abft_dbl(xl), lo);
        hi = fma(w, abft_dbl(xh), hi);
    }
    __device__ __forceinline__ void value(u64& s0, u64& s1) const {
        const u64 l = __double2ull_rn(lo), h = __double2ull_rn(hi);
        s0 = l + (h << 32);
        s1 = (h >> 32) + (s0 < l);
    }
};

struct AbftOutF {
    double p[2][6] = {};
    // om = omega0 | omega1 << 32 (30-bit halves), x canonical (< 2^60); set = 0 / 1 alternates so each accumulator sees 8 terms
    __device__ __forceinline__ void add(int set, u64 om, u64 x) {
        u32 o0, o1;
        unpack64(om, o0, o1);
        const double a0 = abft_dbl(o0), a1 = abft_dbl(o1);
        const double x0 = abft_dbl((u32)x & 0xfffffu), x1 = abft_dbl((u32)(x >> 20) & 0xfffffu), x2 = abft_dbl((u32)(x >> 40));
        p[set][0] = fma(a0, x0, p[set][0]); p[set][1] = fma(a0, x1, p[set][1]); p[set][2] = fma(a0, x2, p[set][2]);
        p[set][3] = fma(a1, x0, p[set][3]); p[set][4] = fma(a1, x1, p[set][4]); p[set][5] = fma(a1, x2, p[set][5]);
    }
    // sum_{i,j} P_ij 2^(30 i + 20 j)  < 2^125
    __device__ __forceinline__ void value(u64& s0, u64& s1) const {
        unsigned __int128 t = 0;
#pragma unroll
        for (int i = 0; i < 2; i++)
#pragma unroll
            for (int j = 0; j < 3; j++)
                t += (unsigned __int128)(__double2ull_rn(p[0][3 * i + j]) + __double2ull_rn(p[1][3 * i + j])) << (30 * i + 20 * j);
        s0 = (u64)t;
        s1 = (u64)(t >> 64);
    }
};

// counters: [limb][side][8] u64, side 0 = input sum, 1 = output sum.  The 128-bit per-thread sum
// (s1:s0) is committed unreduced as NDIG 26-bit digits: one REDUX per digit per warp and one atomic per
// digit per CTA (-DB200_ABFT_WARP_COMMIT: per warp, the r1c form; 17.9 % -> 16.6 % overhead at C2's shape);
// k_abft_finalize recombines the digit sums and reduces mod q once per limb.  CTA-uniform call sites only.
constexpr int kAbftStride = 8;
constexpr int kAbftDigit = 26;       // 32 lanes x 26-bit digits sum to < 2^31: one 32-bit REDUX per digit
// grp / n_grp: the CTA's warps may belong to n_grp different limbs (the row passes process several polynomials side by
// side); the warps [grp * wpg, (grp + 1) * wpg) commit to `limb` of their own group.  Call sites must be uniform per group.
template <int NDIG>
__device__ __forceinline__ void abft_commit(u64* counters, u32 limb, int side, u64 s0, u64 s1, int grp = 0, int n_grp = 1) {
    constexpr u32 mask = (1u << kAbftDigit) - 1u;
#ifndef B200_ABFT_WARP_COMMIT
    __shared__ u32 red[32][NDIG];        // per-warp digit sums (< 2^31 each); every thread of the CTA calls this
    __shared__ u32 glimb[4];
    const u32 wpg = (blockDim.x >> 5) / n_grp;                 // warps per group
    if ((threadIdx.x & 31) == 0 && (threadIdx.x >> 5) % wpg == 0) glimb[grp] = limb;
#else
    u64* c = counters + ((size_t)limb * 2 + side) * kAbftStride;
#endif
#ifdef B200_ABL_ABFT_NOCOMMIT        // timing ablation (profiles/microbench/runexp.sh): sums stay live, no REDUX / RED
    if ((s0 ^ s1) != 0x9E3779B97F4A7C15ull) return;
#endif
#pragma unroll
    for (int d = 0; d < NDIG; d++) {
        const int sh = kAbftDigit * d;
        u32 dig;
        if (sh + kAbftDigit <= 64) dig = (u32)(s0 >> sh) & mask;
        else if (sh >= 64) dig = (u32)(s1 >> (sh - 64)) & mask;
        else dig = (u32)((s0 >> sh) | (s1 << (64 - sh))) & mask;
        const u32 sum = __reduce_add_sync(0xffffffffu, dig);
#ifndef B200_ABFT_WARP_COMMIT
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5][d] = sum;
#else
        // plain RED from lane 0 (inline PTX: atomicAdd here makes nvcc emit its warp-aggregation sequence per digit)
        if ((threadIdx.x & 31) == 0) asm volatile("red.global.add.u64 [%0], %1;" ::"l"(c + d), "l"((u64)sum) : "memory");
#endif
    }
#ifndef B200_ABFT_WARP_COMMIT
    // one RED per digit per group instead of per warp: the counters of a limb are one hot L2 line
    __syncthreads();
    if (threadIdx.x < (u32)(NDIG * n_grp)) {
        const u32 g = threadIdx.x / NDIG, d = threadIdx.x % NDIG;
        u64 tot = 0;
        for (u32 w = g * wpg; w < (g + 1) * wpg; w++) tot += red[w][d];
        u64* c = counters + ((size_t)glimb[g] * 2 + side) * kAbftStride;
        asm volatile("red.global.add.u64 [%0], %1;" ::"l"(c + d), "l"(tot) : "memory");
    }
#endif
}

// generate_weights(N): w_i = (i mod p) + (i div p) + 2, p = 2^pl
__device__ __forceinline__ u32 abft_w(u32 i, int pl) { return (i & ((1u << pl) - 1)) + (i >> pl) + 2; }
