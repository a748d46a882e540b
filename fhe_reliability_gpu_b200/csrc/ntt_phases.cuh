// Register-resident radix-2^LR butterfly phases of the negacyclic NTT.
//
// Indexing (DESIGN.md "twiddle indexing"): SEAL's forward stage with m groups
// reads root_powers[m + i] for group i.  For a sub-transform over a local index
// x of 2^LX points whose global group prefix is P (P = 1 for the column pass,
// P = n_cols_len + r for row r of the row pass), local stage k uses
// tw[(P << k) + (x >> (LX - k))].  A register phase that holds the R = 2^LR
// elements e of x = x_hi*2^(LX-k0) + e*2^(LX-k0-LR) + x_lo therefore sees, at
// its sub-stage t and register group g = e >> (LR - t),
//     tw[(Pp << t) + g],   Pp = (P << k0) + x_hi
// i.e. the same rule with P replaced by Pp.  The inverse (Gentleman-Sande)
// walks the same indices backwards with the inverse table.
#pragma once
#include "modarith.cuh"

// Forward Cooley-Tukey phase.  In: any 64-bit words (congruent to the wanted
// residues).  Out: values < (3 + 2*LR) q  (<= 13q for LR = 5).
template <int LR, bool BIG>
__device__ __forceinline__ void ct_phase(u64 (&v)[1 << LR], const ulonglong2* __restrict__ tw, u32 Pp,
                                         const B200Mod& m) {
    constexpr int R = 1 << LR;
    const u64 q = m.q, two_q = q << 1;
#pragma unroll
    for (int e = 0; e < R / 2; e++) v[e] = reduce_lazy<BIG>(v[e], m);   // X half of sub-stage 0
#pragma unroll
    for (int t = 0; t < LR; t++) {
        const int half = R >> (t + 1);
#pragma unroll
        for (int g = 0; g < (1 << t); g++) {
            const ulonglong2 W = ldg_tw(tw + ((Pp << t) + g));
#pragma unroll
            for (int i = 0; i < half; i++) {
                const int a = g * 2 * half + i, b = a + half;
                const u64 T = shoup_mul(v[b], W.x, W.y, q);
                const u64 X = v[a];
                v[a] = X + T;
                v[b] = X - T + two_q;
            }
        }
    }
}

// Inverse Gentleman-Sande phase.  In: any 64-bit words.  Out: values < 12q
// (LR even) / < 6q (LR odd) unless LAST, in which case the final sub-stage
// (global stage 0) folds N^-1 in and the outputs are canonical [0,q).
template <int LR, bool BIG, bool LAST>
__device__ __forceinline__ void gs_phase(u64 (&v)[1 << LR], const ulonglong2* __restrict__ itw, u32 Pp,
                                         const B200Mod& m) {
    constexpr int R = 1 << LR;
    const u64 q = m.q;
#pragma unroll
    for (int e = 0; e < R; e++) v[e] = reduce_lazy<BIG>(v[e], m);       // all < 3q
    int bound = 3;                                                      // compile-time after unroll
#pragma unroll
    for (int n = 0; n < LR; n++) {
        const int t = LR - 1 - n;
        const int half = R >> (t + 1);
        if (bound > 8) {                       // sums would pass 16q: re-reduce the big (sum) elements
            const int phalf = half >> 1;       // previous sub-stage's half
#pragma unroll
            for (int e = 0; e < R; e++)
                if ((e & phalf) == 0) v[e] = reduce_lazy<BIG>(v[e], m);
            bound = 3;
        }
        const u64 bq = q * (u64)bound;
        const bool fold = LAST && (t == 0);
#pragma unroll
        for (int g = 0; g < (1 << t); g++) {
            ulonglong2 W;
            if (fold) { W.x = m.iw1n; W.y = m.iw1n_s; }
            else W = ldg_tw(itw + ((Pp << t) + g));
#pragma unroll
            for (int i = 0; i < half; i++) {
                const int a = g * 2 * half + i, b = a + half;
                const u64 U = v[a], V = v[b];
                const u64 S = U + V;
                v[b] = shoup_mul(U - V + bq, W.x, W.y, q);
                v[a] = fold ? shoup_mul(S, m.ninv, m.ninv_s, q) : S;
            }
        }
        bound *= 2;
    }
    if (LAST) {
#pragma unroll
        for (int e = 0; e < R; e++) v[e] = canon2(v[e], q);
    }
}
