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
//
// A phase fetches its twiddles through a small functor so the same butterfly
// code runs from the SEAL-layout table in global memory (column passes) and
// from the TMA-staged shared-memory tile of the row passes.
#pragma once
#include "modarith.cuh"

// tw[(Pp << t) + g] from the SEAL-layout (bit-reversed) table in global memory
struct TwGlobal {
    const ulonglong2* __restrict__ tw;
    u32 Pp;
    __device__ __forceinline__ ulonglong2 get(int t, int g) const { return __ldg(tw + ((Pp << t) + g)); }
};
// the same indexing from a copy of the table's first entries staged in shared memory (column passes)
struct TwShared {
    const ulonglong2* tw;
    u32 Pp;
#ifdef B200_ABL_TWREG
    __device__ __forceinline__ ulonglong2 get(int t, int g) const { return tw[Pp]; }
#else
    __device__ __forceinline__ ulonglong2 get(int t, int g) const { return tw[(Pp << t) + g]; }
#endif
};
// entry j = (1 << t) - 1 + g of a per-thread (or per-row) list with a fixed stride (shared memory)
template <int STRIDE>
struct TwList {
    const ulonglong2* base;
#ifdef B200_ABL_TWREG
    __device__ __forceinline__ ulonglong2 get(int t, int g) const { return base[0]; }
#else
    __device__ __forceinline__ ulonglong2 get(int t, int g) const { return base[((1 << t) - 1 + g) * STRIDE]; }
#endif
};

// ---- fault-injection hook (protected path only) ---------------------------------------------
// Mirrors the per-operation injection of rfhe_framewk/src/ntt_test/relia_ntt_sim.cpp:169-206:
// the fault hits v = w * A[k + half] (the butterfly's product) of one chosen operation.  An
// entry names the global stage and the coefficient index of the butterfly's upper ("a") input.
struct InjEntry {
    u32 limb;      // limb index inside the call (poly * L + l)
    u32 stage;     // 0 .. log_n - 1, SEAL stage order (m = 2^stage groups)
    u32 xa;        // coefficient index of the a-input: 2*i*gap + j for group i, offset j
    u32 kind;      // 1: v = (v ^ payload) mod q (SBF/DBF, :118-127)   2: v = payload mod q (MLF, :128-130)
    u64 payload;
};
struct NoInject {
    static constexpr bool on = false;
    __device__ __forceinline__ u64 apply(int, int, u64 T) const { return T; }
};
__device__ __forceinline__ u64 inj_corrupt(u64 T, const InjEntry& e, u64 q) {
    u64 v = T % q;
    return e.kind == 1 ? (v ^ e.payload) % q : e.payload % q;
}
// element e of the phase sits at coefficient index base + (e << shift); phase sub-stage t is global stage stage0 + t
struct Inject {
    static constexpr bool on = true;
    const InjEntry* e;
    u32 n, limb, base, shift, stage0;
    u64 q;
    __device__ __forceinline__ u64 apply(int t, int a, u64 T) const {
        for (u32 i = 0; i < n; i++)
            if (e[i].limb == limb && e[i].stage == stage0 + (u32)t && e[i].xa == base + ((u32)a << shift)) T = inj_corrupt(T, e[i], q);
        return T;
    }
};

// Forward Cooley-Tukey phase.  In: any 64-bit words (congruent to the wanted
// residues).  Out: values < 16q.  Lazy products are < 3q, so the X path grows by 3q per
// sub-stage from reduce_lazy's bound B0 <= 3: 15q after four sub-stages; a fifth
// sub-stage (LR = 5, the N = 2^17 column pass) re-reduces its X inputs first.
template <int LR, int BIG, class TW, class INJ = NoInject>
__device__ __forceinline__ void ct_phase(u64 (&v)[1 << LR], const TW& tw, const B200Mod& m, const INJ& inj = INJ()) {
    constexpr int R = 1 << LR;
#ifdef B200_ABL_NOCOMPUTE     // timing ablation: memory side only
    if (m.q != 12345) return;
#endif
    const u64 q = m.q, q3 = m.q3, nq = 0 - q;
    const u32 sh = m.bits - 32;            // NEAR only
#pragma unroll
    for (int e = 0; e < R / 2; e++) v[e] = reduce_lazy<BIG>(v[e], m);   // X half of sub-stage 0
#pragma unroll
    for (int t = 0; t < LR; t++) {
        const int half = R >> (t + 1);
        if (t == 4) {                          // X inputs would pass 16q - 3q: bring them back below 3q
#pragma unroll
            for (int e = 0; e < R; e++)
                if ((e & half) == 0) v[e] = reduce_lazy<BIG>(v[e], m);
        }
#pragma unroll
        for (int g = 0; g < (1 << t); g++) {
            const ulonglong2 W = tw.get(t, g);
#pragma unroll
            for (int i = 0; i < half; i++) {
                const int a = g * 2 * half + i, b = a + half;
                u64 T = shoup_mul_lazy<BIG>(v[b], W.x, W.y, nq, sh);
                if (INJ::on) T = inj.apply(t, a, T);
                const u64 X = v[a];
                v[a] = X + T;
                v[b] = X - T + q3;
            }
        }
    }
}

// ---- lazy-range plan of a Gentleman-Sande phase (compile time) -----------------------------------------------------
// Values ride as multiples of q below 16q (q < 2^60).  A GS butterfly (U, V) -> (U + V, (U - V + c q) w) needs
// bound(U) + bound(V) <= 16 and bound(U) + c <= 16 with c q >= V; products come back below 3q, sums add their bounds.
// The r1 code tracked ONE worst-case bound for all elements and re-reduced half the vector when it passed 8: 24
// reductions per radix-16 phase.  Tracking the bound of every element (the butterfly network is fixed, so this is pure
// compile-time bookkeeping) shows that only the few elements on an all-sums path ever get close to 16q: 19 reductions
// when the inputs are unknown words (16 of them the input reduction itself), 3-9 when the producer guarantees a bound.
// IN_B: bound of every input in units of q (0: arbitrary 64-bit words -> all are reduced first).
// OUT_T: outputs whose bound exceeds OUT_T are reduced at the end (0: leave them), so the consumer can start from OUT_T.
struct GsPlan {
    unsigned red[6];             // [n]: elements reduced before sub-stage n (bit e); [LR]: reduced at the end
    unsigned char c[5][16];      // [n][butterfly]: multiple of q added to U - V
    unsigned char out[32];       // output bounds (after the end reductions)
    unsigned char out_max;
};
__host__ __device__ constexpr int gs_const_ge(int b) {
    const int cs[8] = {1, 2, 3, 4, 6, 8, 12, 16};
    for (int i = 0; i < 8; i++)
        if (cs[i] >= b) return cs[i];
    return 0;
}
template <int LR, int LAZYB, int IN_B, int OUT_T, bool LAST>
__host__ __device__ constexpr GsPlan gs_make_plan() {
    constexpr int R = 1 << LR;
    GsPlan p{};
    int bnd[32] = {};
    for (int e = 0; e < R; e++) bnd[e] = IN_B ? IN_B : LAZYB;
    if (!IN_B) p.red[0] = R == 32 ? 0xffffffffu : ((1u << R) - 1u);
    for (int n = 0; n < LR; n++) {
        const int t = LR - 1 - n, half = R >> (t + 1);
        int bf = 0;
        for (int g = 0; g < (1 << t); g++)
            for (int i = 0; i < half; i++, bf++) {
                const int a = g * 2 * half + i, b = a + half;
                int ba = bnd[a], bb = bnd[b], c = gs_const_ge(bb);
                if (!(ba + bb <= 16 && c && ba + c <= 16)) {
                    // cheapest fix first: V alone, U alone, both
                    const int cb = gs_const_ge(LAZYB);
                    if (ba + LAZYB <= 16 && ba + cb <= 16) { p.red[n] |= 1u << b; bb = LAZYB; }
                    else if (LAZYB + bb <= 16 && c && LAZYB + c <= 16) { p.red[n] |= 1u << a; ba = LAZYB; }
                    else { p.red[n] |= (1u << a) | (1u << b); ba = bb = LAZYB; }
                    c = gs_const_ge(bb);
                }
                p.c[n][bf] = (unsigned char)c;
                bnd[a] = (LAST && t == 0) ? 3 : ba + bb;
                bnd[b] = 3;
            }
    }
    for (int e = 0; e < R; e++) {
        if (OUT_T && bnd[e] > OUT_T) { p.red[LR] |= 1u << e; bnd[e] = LAZYB; }
        p.out[e] = (unsigned char)bnd[e];
        if (bnd[e] > p.out_max) p.out_max = (unsigned char)bnd[e];
    }
    return p;
}

// Inverse Gentleman-Sande phase.  In: words below IN_B q (IN_B = 0: any 64-bit words).  Out: lazily reduced values
// below gs_out_bound<...>() q unless LAST, in which case the final sub-stage (global stage 0) folds N^-1 in and the
// outputs are canonical [0,q).
template <int LR, int BIG, bool LAST, int IN_B, int OUT_T>
__host__ __device__ constexpr int gs_out_bound() { return gs_make_plan<LR, LazyBound<BIG>::value, IN_B, OUT_T, LAST>().out_max; }

template <int LR, int BIG, bool LAST, int IN_B = 0, int OUT_T = 0, class TW>
__device__ __forceinline__ void gs_phase(u64 (&v)[1 << LR], const TW& tw, const B200Mod& m) {
    constexpr int R = 1 << LR;
#ifdef B200_ABL_NOCOMPUTE
    if (m.q != 12345) return;
#endif
    constexpr GsPlan P = gs_make_plan<LR, LazyBound<BIG>::value, IN_B, OUT_T, LAST>();
    const u64 q = m.q, nq = 0 - q;
    const u32 sh = m.bits - 32;            // NEAR only
#pragma unroll
    for (int n = 0; n < LR; n++) {
        const int t = LR - 1 - n;
        const int half = R >> (t + 1);
#pragma unroll
        for (int e = 0; e < R; e++)
            if ((P.red[n] >> e) & 1u) v[e] = reduce_lazy<BIG>(v[e], m);
        const bool fold = LAST && (t == 0);
#pragma unroll
        for (int g = 0; g < (1 << t); g++) {
            ulonglong2 W;
            if (fold) { W.x = m.iw1n; W.y = m.iw1n_s; }
            else W = tw.get(t, g);
#pragma unroll
            for (int i = 0; i < half; i++) {
                const int a = g * 2 * half + i, b = a + half;
                // c q with c in {1,2,3,4,6,8,12,16}: shifts of q or of the memory-resident 3q (no multiplier-pipe rematerialisation)
                const int c = P.c[n][g * half + i];
                const u64 bq = (c % 3 == 0) ? (m.q3 << (31 - __builtin_clz(c / 3))) : (q << (31 - __builtin_clz(c)));
                const u64 U = v[a], V = v[b];
                const u64 S = U + V;
                v[b] = shoup_mul_lazy<BIG>(U - V + bq, W.x, W.y, nq, sh);
                v[a] = fold ? shoup_mul_lazy<BIG>(S, m.ninv, m.ninv_s, nq, sh) : S;
            }
        }
    }
#pragma unroll
    for (int e = 0; e < R; e++) {
        if ((P.red[LR] >> e) & 1u) v[e] = reduce_lazy<BIG>(v[e], m);
        if (LAST) v[e] = canon3(v[e], q);
    }
}
