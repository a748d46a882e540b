// Batched negacyclic NTT / INTT over RNS limbs, hand-written for sm_100a.
//
// Replaces phantom-fhe's nwt_2d_radix8_forward_inplace / backward (call sites
// reliability_test/ntt_test.cu:95,144; definition absent from the reference
// tree) with SEAL ntt_negacyclic_harvey semantics: forward natural -> bit-
// reversed, inverse bit-reversed -> natural scaled by N^-1, canonical outputs.
//
// Structure (N = CL x 256, 2^12 <= N <= 2^17): the limb is a CL x 256 matrix.
//   forward : column pass (first log2 CL stages, twiddles shared by all columns)
//             then row pass (last 8 stages, per-row twiddles, contiguous rows);
//   inverse : row pass then column pass (N^-1 folded into the last stage).
// Each pass keeps 16/32 coefficients per thread in registers for a radix-16/32
// phase, exchanges once through shared memory, and runs the second phase.
// Intermediates are stored lazily (< 16q); only final outputs are canonical.
#include "ntt_phases.cuh"
#include "ctx.hpp"
#include "abft.cuh"
#include <atomic>
#include <cstdlib>

namespace b200 {

constexpr int kRowLog = 8;          // row length 256
constexpr int kRowLen = 1 << kRowLog;
constexpr int kTile = 16;           // columns per column-pass CTA / rows per row-pass CTA
// Thread groups (of kTile threads) per column-pass CTA.  The two register phases of a column need 2^LB and 2^LA groups.
// Up to LC = 8 the CTA has max(2^LA, 2^LB) groups (equal there, or the small phase idles a few).  For LC = 9 (radix 32 then
// radix 16) that left half of the 512 threads idle during the radix-32 phase while their 128 registers kept a second CTA
// off the SM: the CTA now has min(2^LA, 2^LB) = 16 groups, every thread runs the radix-32 phase and loops twice over the
// radix-16 phase, and two CTAs fit (N = 2^17: 1.84 -> see profiles/microbench/ntt_size_sweep.py).
__host__ __device__ constexpr int col_groups(int LC, int LA) {
    const int LB = LC - LA, hi = LA > LB ? LA : LB, lo = LA > LB ? LB : LA;
    return 1 << (LC >= 9 ? lo : hi);
}



// ---------------------------------------------------------------- TMA helpers
// 1-D bulk async copy global -> shared, completion counted on an mbarrier.
__device__ __forceinline__ u32 smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u64* bar, u32 count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(u64* bar, u32 bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, u32 bytes, u64* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(u64* bar, u32 parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

// ---------------------------------------------------------------- forward, columns
// PROT: the protected transform (config C4) -- accumulates the input side of the
// ABFT checksum while the coefficients are loaded and honours the injection plan.
// resident CTAs per SM the column kernels are compiled for (register cap): 3 x 256 threads (85 registers); a cap of 4 (64
// registers) spills and measured no faster; 512-row columns (LC = 9) fit 2.  The inverse kernel at LC = 8 spills 16 bytes under the
// cap of 3; with 2 CTAs (106 registers, no spill) it measured 111 us against 105 us per 288-limb batch: the cap of 3 stays
__host__ __device__ constexpr int col_min_ctas(int LC, bool inverse) { return LC >= 9 ? 2 : 3; }
// MAP: the launch's entries come from the scatter list p.map (data limb, modulus) instead of (poly L + y, start_mod + y)
template <int LC, int LA, int BIG, bool PROT, bool INJ, bool MAP = false>
__global__ void __launch_bounds__(kTile * col_groups(LC, LA), INJ ? 1 : col_min_ctas(LC, false))
k_fwd_col(PassArgs p) {
    constexpr int LB = LC - LA, RA = 1 << LA, RB = 1 << LB, NG = col_groups(LC, LA);
    static_assert(NG >= RB, "the first phase runs in one sweep");
    extern __shared__ u64 sm[];
    // grid: x = (tile, poly) with poly fastest, y = limb of the polynomial.  CTAs that share a modulus (and so a
    // twiddle tile) are dispatched back to back: the tile is read from HBM once and hit in L2 by the other polys.
    const u32 poly = blockIdx.x % p.n_poly, tile = blockIdx.x / p.n_poly;
    const u32 limb = poly * p.L + blockIdx.y;
    u32 mi = p.start_mod + blockIdx.y, dl = limb;
    if constexpr (MAP) { const uint2 e = p.map[blockIdx.y]; dl = e.x; mi = e.y; }
    const B200Mod m = p.mods[mi];
    const ulonglong2* tw = p.tw + (size_t)mi * p.N;
    u64* base = p.data + (size_t)dl * p.N + tile * kTile;
    const int c = threadIdx.x & 15, g = threadIdx.x >> 4;
    const u32 col = tile * kTile + c;
    u64 ck0 = 0, ck1 = 0;
    // the 2^LC - 1 twiddles of the column stages are shared by every thread of the CTA: stage them once
    ulonglong2* stw = reinterpret_cast<ulonglong2*>(sm + (LB > 0 ? (kTile << LC) : 0));
    u64 v[RA];
    if (g < RB) {
#pragma unroll
        for (int e = 0; e < RA; e++) v[e] = base[(size_t)(e * RB + g) * kRowLen + c];
    }
    for (int i = threadIdx.x; i < (1 << LC); i += blockDim.x) stw[i] = __ldg(tw + i);
    __syncthreads();
    if (g < RB) {
        if constexpr (PROT) {
            constexpr int pl = (LC + kRowLog + 1) / 2;
            static_assert(pl <= LB + kRowLog, "the weights of a thread's elements are affine in e");
#if defined(B200_ABFT_INT) || defined(B200_ABFT_COL_INT)                 // the r1 form: weighted products on the integer-multiply pipe
            AbftIn acc;
#pragma unroll
            for (int e = 0; e < RA; e++) acc.add(abft_w((u32)(e * RB + g) * kRowLen + col, pl), v[e]);
#else
            // element e sits at index i0 + e (RB 256), and 2^pl divides RB 256: w_e = w_0 + e D  (abft_w = i mod 2^pl + i div 2^pl + 2)
            AbftInF acc;
            const u32 i0 = (u32)g * kRowLen + col;
            double wd = (double)abft_w(i0, pl);
            const double dd = (double)((RB * kRowLen) >> pl);
#pragma unroll
            for (int e = 0; e < RA; e++) { acc.add(wd, v[e]); wd += dd; asm("" : "+d"(wd)); }   // keep the DADD chain (else 16 I2F.F64)
#endif
            acc.value(ck0, ck1);
        }
        if constexpr (INJ)
            ct_phase<LA, BIG>(v, TwShared{stw, 1u}, m,
                              Inject{p.inj, p.n_inj, p.limb0 + limb, (u32)g * kRowLen + col, (u32)(LB + kRowLog), 0u, m.q});
        else
            ct_phase<LA, BIG>(v, TwShared{stw, 1u}, m);
        if (LB > 0) {
#pragma unroll
            for (int e = 0; e < RA; e++) sm[(e * RB + g) * kTile + c] = v[e];
        } else {
#pragma unroll
            for (int e = 0; e < RA; e++) base[(size_t)(e * RB + g) * kRowLen + c] = v[e];
        }
    }
    if constexpr (PROT) abft_commit<4>(p.abft, p.limb0 + limb, 0, ck0, ck1);
    if constexpr (LB > 0) {
        __syncthreads();
#pragma unroll 1
        for (int gb = g; gb < RA; gb += NG) {
            u64 w[RB];
#pragma unroll
            for (int e = 0; e < RB; e++) w[e] = sm[(gb * RB + e) * kTile + c];
            if constexpr (INJ)
                ct_phase<LB, BIG>(w, TwShared{stw, (1u << LA) + gb}, m,
                                  Inject{p.inj, p.n_inj, p.limb0 + limb, (u32)(gb * RB) * kRowLen + col, (u32)kRowLog, (u32)LA, m.q});
            else
                ct_phase<LB, BIG>(w, TwShared{stw, (1u << LA) + gb}, m);
#pragma unroll
            for (int e = 0; e < RB; e++) base[(size_t)(gb * RB + e) * kRowLen + c] = w[e];
        }
    }
}

// Lazy-range contract between the phases of the inverse transform (ntt_phases.cuh GsPlan): every phase reduces its
// outputs that exceed kGsT q before handing them on, so the next phase (after the shared-memory exchange, or in the
// column kernel after the trip through HBM) starts from a known bound instead of reducing everything again.
constexpr int kGsT = 3;
template <int BIG> __host__ __device__ constexpr int inv_row_b1() { return gs_out_bound<4, BIG, false, 0, kGsT>(); }              // after the row pass's first phase
template <int BIG> __host__ __device__ constexpr int inv_row_b2() { return gs_out_bound<4, BIG, false, inv_row_b1<BIG>(), kGsT>(); }   // what the row pass stores

// ---------------------------------------------------------------- row passes
// One CTA = kRowsPerCta (8) rows of 256 coefficients of one limb, one row per half-warp, 16
// coefficients per thread.  The 255 twiddles each row needs (x 8 rows = one
// 32 KiB tile of the row table, laid out so that every shared-memory read is a
// conflict-free 128-bit access) arrive by ONE TMA bulk copy that overlaps the
// coefficient loads; the two radix-16 register phases exchange through a
// 16-byte-chunk XOR-swizzled shared tile.
// NP polynomials that share the modulus (and so the twiddle tile) are processed side by side by one CTA (NP = 2 when the
// batch has at least two): the 32 KiB tile is fetched once per NP x 16 KiB of coefficients instead of once per 16 KiB
// (the row passes move twice as many twiddle bytes as data bytes through L2 otherwise), and the twiddle footprint per
// warp halves, which lifts the shared-memory occupancy limit from 4 x 4 to 3 x 8 warps per SM.
constexpr int kRowSlice = kRowsPerCta * 16;         // threads working on one polynomial's rows
template <int NP>
struct RowSmem {
    ulonglong2 twA[kRowsPerCta * 16];    // [row][15 + pad]: stages 0..3 of the row (broadcast reads)
    ulonglong2 twB[15 * kRowSlice];      // [j][thread]:     stages 4..7, one list per thread of a slice
    u64 x[NP * kRowsPerCta * kRowLen];   // exchange tiles, one per polynomial
    u64 bar;
};
constexpr u32 kRowTileBytes = sizeof(ulonglong2) * kRowTileEntries;
template <int NP> constexpr int row_ctas_per_sm() {
    return (220 * 1024) / (int)sizeof(RowSmem<NP>) > 16 / NP ? 16 / NP : (220 * 1024) / (int)sizeof(RowSmem<NP>);
}

template <int BIG, bool PROT, bool INJ, int NP, bool MAP = false>
__global__ void __launch_bounds__(kRowSlice * NP, row_ctas_per_sm<NP>()) k_fwd_row(PassArgs p) {
    extern __shared__ __align__(128) unsigned char smraw[];
    RowSmem<NP>& s = *reinterpret_cast<RowSmem<NP>*>(smraw);
    // grid: x = (tile, poly group) with the group fastest, y = limb of the polynomial.  CTAs that share a modulus (and
    // so a twiddle tile) are dispatched back to back: the tile is read from HBM once and hit in L2 by the other groups.
    const u32 ng = p.n_poly / NP, grp = blockIdx.x % ng, tile = blockIdx.x / ng;
    const int tslice = threadIdx.x % kRowSlice, pslot = threadIdx.x / kRowSlice;
    const u32 poly = p.poly0 + grp * NP + pslot;
    const u32 limb = poly * p.L + blockIdx.y;
    u32 mi = p.start_mod + blockIdx.y, dl = limb;
    if constexpr (MAP) { const uint2 e = p.map[blockIdx.y]; dl = e.x; mi = e.y; }
    const int cl = tslice & 15, rl = tslice >> 4;
#ifndef B200_ABL_NOTMA
    if (threadIdx.x == 0) {
        mbar_init(&s.bar, 1);
        mbar_expect_tx(&s.bar, kRowTileBytes);
        bulk_g2s(s.twA, p.rtw + (size_t)mi * p.N + (size_t)tile * (kRowTileBytes / 16), kRowTileBytes, &s.bar);
    }
#endif
    const B200Mod m = p.mods[mi];
    u64* rowp = p.data + (size_t)dl * p.N + (size_t)(tile * kRowsPerCta + rl) * kRowLen;
    u64 v[16];
#pragma unroll
#ifdef B200_ABL_NOLOAD
    for (int e = 0; e < 16; e++) v[e] = (u64)threadIdx.x * 0x9E3779B97F4A7C15ull + e + (u64)rowp;
#else
    for (int e = 0; e < 16; e++) v[e] = rowp[e * 16 + cl];
#endif
#ifndef B200_ABL_NOTMA
    __syncthreads();                     // barrier initialised before anyone polls it
    mbar_wait(&s.bar, 0);
#endif
    const u32 x0 = (tile * kRowsPerCta + rl) * kRowLen;     // first coefficient of the row
    const u32 lc = 31 - __clz(p.N >> kRowLog);                     // stages done by the column pass
    if constexpr (INJ) ct_phase<4, BIG>(v, TwList<1>{s.twA + rl * 16}, m, Inject{p.inj, p.n_inj, p.limb0 + limb, x0 + cl, 4u, lc, m.q});
    else ct_phase<4, BIG>(v, TwList<1>{s.twA + rl * 16}, m);
    // exchange: thread cl holds positions e*16+cl, wants cl*16+e
    unsigned char* xrow = reinterpret_cast<unsigned char*>(s.x) + (pslot * kRowsPerCta + rl) * (kRowLen * 8);
    {
        const u32 b0 = cl * 8;
#pragma unroll
        for (int e = 0; e < 16; e++) *reinterpret_cast<u64*>(xrow + e * 128 + (b0 ^ ((e & 7) << 4))) = v[e];
    }
    __syncwarp();                        // a row lives inside one half-warp
    {
        const u32 a0 = cl * 128 + ((cl & 7) << 4);
#pragma unroll
        for (int c = 0; c < 8; c++) {
            const ulonglong2 t = *reinterpret_cast<const ulonglong2*>(xrow + (a0 ^ (c << 4)));
            v[2 * c] = t.x;
            v[2 * c + 1] = t.y;
        }
    }
    if constexpr (PROT) {
        // the exchange line this thread just read is free: stream its 16 output weights into it
        // (cp.async, same chunk swizzle) while the second register phase runs
        __syncwarp();
        const u32 a0 = cl * 128 + ((cl & 7) << 4);
        const unsigned char* om = reinterpret_cast<const unsigned char*>(p.omega + (size_t)mi * p.N + x0 + cl * 16);
#pragma unroll
#ifndef B200_ABL_ABFT_NOOMEGA        // timing ablation: the weights are whatever the exchange tile holds
        for (int c = 0; c < 8; c++)
#else
        for (int c = 0; c < 0; c++)
#endif
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(xrow + (a0 ^ (c << 4)))), "l"(om + c * 16) : "memory");
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    if constexpr (INJ)
        ct_phase<4, BIG>(v, TwList<kRowSlice>{s.twB + tslice}, m, Inject{p.inj, p.n_inj, p.limb0 + limb, x0 + cl * 16, 0u, lc + 4, m.q});
    else
        ct_phase<4, BIG>(v, TwList<kRowSlice>{s.twB + tslice}, m);
#pragma unroll
    for (int e = 0; e < 16; e++) v[e] = BIG ? canon2(reduce_lazy<BIG>(v[e], m), m.q) : canon3(reduce_lazy<BIG>(v[e], m), m.q);
    ulonglong2* out = reinterpret_cast<ulonglong2*>(rowp + cl * 16);
#pragma unroll
#ifdef B200_ABL_NOSTORE
    { u64 acc = 0; for (int e = 0; e < 16; e++) acc ^= v[e]; if (acc == 0x1234567) out[0] = make_ulonglong2(acc, acc); }
#else
    for (int e = 0; e < 8; e++) out[e] = make_ulonglong2(v[2 * e], v[2 * e + 1]);
#endif
    if constexpr (PROT) {
        // output side of the checksum, on the values just stored
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        const u32 a0 = cl * 128 + ((cl & 7) << 4);
#ifdef B200_ABFT_INT
        AbftOut acc;
#pragma unroll
        for (int c = 0; c < 8; c++) {
            const ulonglong2 o = *reinterpret_cast<const ulonglong2*>(xrow + (a0 ^ (c << 4)));
            acc.add(o.x, v[2 * c]);
            acc.add(o.y, v[2 * c + 1]);
        }
#else
        AbftOutF acc;
#pragma unroll
        for (int c = 0; c < 8; c++) {
            const ulonglong2 o = *reinterpret_cast<const ulonglong2*>(xrow + (a0 ^ (c << 4)));
            acc.add(0, o.x, v[2 * c]);
            acc.add(1, o.y, v[2 * c + 1]);
        }
#endif
        u64 s0, s1;
        acc.value(s0, s1);
        abft_commit<5>(p.abft, p.limb0 + limb, 1, s0, s1, pslot, NP);
    }
}

// PROT: the protected inverse transform -- the identity of the forward one read backwards: the inputs here are the
// evaluations (SEAL order), so this pass accumulates sum_k omega_k a^_k on the words it loads (dense weights: FP64 pipe),
// and the column pass accumulates sum_i w_i a_i on the coefficients it stores.
// OOP: out of place -- this (first) pass reads its rows from p.src instead of p.data
template <int BIG, int NP, bool PROT, bool OOP = false>
__global__ void __launch_bounds__(kRowSlice * NP, row_ctas_per_sm<NP>()) k_inv_row(PassArgs p) {
    extern __shared__ __align__(128) unsigned char smraw[];
    RowSmem<NP>& s = *reinterpret_cast<RowSmem<NP>*>(smraw);
    // grid as in k_fwd_row: x = (tile, poly group) with the group fastest, y = limb
    const u32 ng = p.n_poly / NP, grp = blockIdx.x % ng, tile = blockIdx.x / ng;
    const int tslice = threadIdx.x % kRowSlice, pslot = threadIdx.x / kRowSlice;
    const u32 poly = p.poly0 + grp * NP + pslot;
    const u32 limb = poly * p.L + blockIdx.y, mi = p.start_mod + blockIdx.y;
    const int cl = tslice & 15, rl = tslice >> 4;
    if (threadIdx.x == 0) {
        mbar_init(&s.bar, 1);
        mbar_expect_tx(&s.bar, kRowTileBytes);
        bulk_g2s(s.twA, p.rtw + (size_t)mi * p.N + (size_t)tile * (kRowTileBytes / 16), kRowTileBytes, &s.bar);
    }
    const B200Mod m = p.mods[mi];
    u64* rowp = p.data + (size_t)limb * p.N + (size_t)(tile * kRowsPerCta + rl) * kRowLen;
    u64 v[16];
    u64 ck0 = 0, ck1 = 0;
    const ulonglong2* in = reinterpret_cast<const ulonglong2*>((OOP ? p.src + (rowp - p.data) : rowp) + cl * 16);
#pragma unroll
    for (int e = 0; e < 8; e++) { ulonglong2 x = in[e]; v[2 * e] = x.x; v[2 * e + 1] = x.y; }
    if constexpr (PROT) {
        // every loaded word is taken mod q (the transform's contract), and it is that canonical value the checksum weighs
        const ulonglong2* om = reinterpret_cast<const ulonglong2*>(p.omega + (size_t)mi * p.N + (size_t)(tile * kRowsPerCta + rl) * kRowLen + cl * 16);
        // the weighted products on the integer-multiply pipe (4 IMAD.WIDE per word); -DB200_ABFT_INV_F64 puts them on the FP64 pipe
        // as the forward row pass does -- measured here 34.6 % against 32.0 % overhead (profiles/r3f_inv_prot_{f64,int}.txt): at the START of
        // a kernel the shorter instruction stream wins, at the end (forward) the two were equal
#ifndef B200_ABFT_INV_F64
        AbftOut acc;
#else
        AbftOutF acc;
#endif
#pragma unroll
        for (int e = 0; e < 8; e++) {
            const ulonglong2 o = __ldg(om + e);
            v[2 * e] = BIG ? canon2(reduce_lazy<BIG>(v[2 * e], m), m.q) : canon3(reduce_lazy<BIG>(v[2 * e], m), m.q);
            v[2 * e + 1] = BIG ? canon2(reduce_lazy<BIG>(v[2 * e + 1], m), m.q) : canon3(reduce_lazy<BIG>(v[2 * e + 1], m), m.q);
#ifndef B200_ABFT_INV_F64
            acc.add(o.x, v[2 * e]);
            acc.add(o.y, v[2 * e + 1]);
#else
            acc.add(0, o.x, v[2 * e]);
            acc.add(1, o.y, v[2 * e + 1]);
#endif
        }
        acc.value(ck0, ck1);             // committed at the END of the kernel: the FP64 chain then overlaps the first butterfly phase
    }
    __syncthreads();                     // barrier initialised before anyone polls it
    mbar_wait(&s.bar, 0);
    gs_phase<4, BIG, false, PROT ? 1 : 0, kGsT>(v, TwList<kRowSlice>{s.twB + tslice}, m);
    unsigned char* xrow = reinterpret_cast<unsigned char*>(s.x) + (pslot * kRowsPerCta + rl) * (kRowLen * 8);
    {
        const u32 a0 = cl * 128 + ((cl & 7) << 4);
#pragma unroll
        for (int c = 0; c < 8; c++)
            *reinterpret_cast<ulonglong2*>(xrow + (a0 ^ (c << 4))) = make_ulonglong2(v[2 * c], v[2 * c + 1]);
    }
    __syncwarp();
    {
        const u32 b0 = cl * 8;
#pragma unroll
        for (int e = 0; e < 16; e++) v[e] = *reinterpret_cast<const u64*>(xrow + e * 128 + (b0 ^ ((e & 7) << 4)));
    }
    constexpr int b1 = PROT ? gs_out_bound<4, BIG, false, 1, kGsT>() : inv_row_b1<BIG>();
    static_assert(gs_out_bound<4, BIG, false, b1, kGsT>() <= inv_row_b2<BIG>(), "the column pass assumes inv_row_b2");
    gs_phase<4, BIG, false, b1, kGsT>(v, TwList<1>{s.twA + rl * 16}, m);
#pragma unroll
    for (int e = 0; e < 16; e++) rowp[e * 16 + cl] = v[e];
    if constexpr (PROT) abft_commit<5>(p.abft, p.limb0 + limb, 0, ck0, ck1, pslot, NP);
}

// ---------------------------------------------------------------- inverse, columns
template <int LC, int LA, int BIG, bool PROT>
__global__ void __launch_bounds__(kTile * col_groups(LC, LA), col_min_ctas(LC, true))
k_inv_col(PassArgs p) {
    constexpr int LB = LC - LA, RA = 1 << LA, RB = 1 << LB, NG = col_groups(LC, LA);
    static_assert(NG >= RB, "the last phase runs in one sweep");
    extern __shared__ u64 sm[];
    // grid: x = (tile, poly) with poly fastest, y = limb of the polynomial.  CTAs that share a modulus (and so a
    // twiddle tile) are dispatched back to back: the tile is read from HBM once and hit in L2 by the other polys.
    const u32 poly = blockIdx.x % p.n_poly, tile = blockIdx.x / p.n_poly;
    const u32 limb = poly * p.L + blockIdx.y, mi = p.start_mod + blockIdx.y;
    const B200Mod m = p.mods[mi];
    const ulonglong2* itw = p.tw + (size_t)mi * p.N;
    u64* base = p.data + (size_t)limb * p.N + tile * kTile;
    const int c = threadIdx.x & 15, g = threadIdx.x >> 4;
    ulonglong2* stw = reinterpret_cast<ulonglong2*>(sm + (LB > 0 ? (kTile << LC) : 0));
    u64 ck0 = 0, ck1 = 0;
    for (int i = threadIdx.x; i < (1 << LC); i += blockDim.x) stw[i] = __ldg(itw + i);
    if constexpr (LB == 0) __syncthreads();
    if constexpr (LB > 0) {
        if constexpr (NG >= RA) {
            u64 w[RB];
            if (g < RA) {
#pragma unroll
                for (int e = 0; e < RB; e++) w[e] = base[(size_t)(g * RB + e) * kRowLen + c];
            }
            __syncthreads();
            if (g < RA) {
                gs_phase<LB, BIG, false, inv_row_b2<BIG>(), kGsT>(w, TwShared{stw, (1u << LA) + g}, m);
#pragma unroll
                for (int e = 0; e < RB; e++) sm[(g * RB + e) * kTile + c] = w[e];
            }
        } else {
            __syncthreads();                                   // twiddles staged
#pragma unroll 1
            for (int gb = g; gb < RA; gb += NG) {
                u64 w[RB];
#pragma unroll
                for (int e = 0; e < RB; e++) w[e] = base[(size_t)(gb * RB + e) * kRowLen + c];
                gs_phase<LB, BIG, false, inv_row_b2<BIG>(), kGsT>(w, TwShared{stw, (1u << LA) + gb}, m);
#pragma unroll
                for (int e = 0; e < RB; e++) sm[(gb * RB + e) * kTile + c] = w[e];
            }
        }
        __syncthreads();
    }
    if (g < RB) {
        u64 v[RA];
        if (LB > 0) {
#pragma unroll
            for (int e = 0; e < RA; e++) v[e] = sm[(e * RB + g) * kTile + c];
        } else {
#pragma unroll
            for (int e = 0; e < RA; e++) v[e] = base[(size_t)(e * RB + g) * kRowLen + c];
        }
        gs_phase<LA, BIG, true, (LB > 0 ? gs_out_bound<(LB > 0 ? LB : 1), BIG, false, inv_row_b2<BIG>(), kGsT>() : inv_row_b2<BIG>()), 0>(v, TwShared{stw, 1u}, m);
#pragma unroll
        for (int e = 0; e < RA; e++) base[(size_t)(e * RB + g) * kRowLen + c] = v[e];
        if constexpr (PROT) {
            // coefficient side of the checksum on the canonical values just stored; weights affine in e as in k_fwd_col
            constexpr int pl = (LC + kRowLog + 1) / 2;
            static_assert(pl <= LB + kRowLog, "the weights of a thread's elements are affine in e");
            AbftInF acc;
            double wd = (double)abft_w((u32)g * kRowLen + tile * kTile + c, pl);
            const double dd = (double)((RB * kRowLen) >> pl);
#pragma unroll
            for (int e = 0; e < RA; e++) { acc.add(wd, v[e]); wd += dd; asm("" : "+d"(wd)); }   // keep the DADD chain (else 16 I2F.F64)
            acc.value(ck0, ck1);
        }
    }
    if constexpr (PROT) abft_commit<4>(p.abft, p.limb0 + limb, 1, ck0, ck1);
}

// ---------------------------------------------------------------- small N (2 <= N <= 2^11)
// One CTA per limb, whole limb in shared memory, radix-2 stages, fully reduced.
// Serves the prototype-sized cases (N = 16 .. 2048) of the Python bindings.
template <bool INVERSE, bool PROT>
__global__ void __launch_bounds__(256) k_small(PassArgs p, int logn) {
    extern __shared__ u64 a[];
    const u32 limb = blockIdx.x, mi = p.start_mod + limb % p.L;
    const B200Mod m = p.mods[mi];
    const ulonglong2* tw = p.tw + (size_t)mi * p.N;
    u64* base = p.data + (size_t)limb * p.N;
    const u32 N = p.N;
    const u64 nq = 0 - m.q;
    AbftIn ain;
    AbftOut aev;
    for (u32 i = threadIdx.x; i < N; i += blockDim.x) {
        const u64 x = base[i];
        if (PROT && !INVERSE) ain.add(abft_w(i, (logn + 1) / 2), x);
        a[i] = canon3(reduce_lazy<false>(x, m), m.q);
        if (PROT && INVERSE) aev.add(p.omega[(size_t)mi * N + i], a[i]);      // evaluation side: sum_k omega_k a^_k
    }
    if (PROT && !INVERSE) { u64 s0, s1; ain.value(s0, s1); abft_commit<4>(p.abft, p.limb0 + limb, 0, s0, s1); }
    if (PROT && INVERSE) { u64 s0, s1; aev.value(s0, s1); abft_commit<5>(p.abft, p.limb0 + limb, 0, s0, s1); }
    __syncthreads();
    if (!INVERSE) {
        for (int s = 0; s < logn; s++) {
            const u32 t = N >> (s + 1), mm = 1u << s;
            for (u32 b = threadIdx.x; b < N / 2; b += blockDim.x) {
                const u32 i = b / t, j = b % t, x = 2 * i * t + j;
                const ulonglong2 W = tw[mm + i];
                u64 T = canon2(shoup_mul(a[x + t], W.x, W.y, nq), m.q);
                if (PROT)
                    for (u32 k = 0; k < p.n_inj; k++)
                        if (p.inj[k].limb == p.limb0 + limb && p.inj[k].stage == (u32)s && p.inj[k].xa == x) T = inj_corrupt(T, p.inj[k], m.q);
                const u64 X = a[x];
                a[x] = canon2(X + T, m.q);
                a[x + t] = canon2(X + m.q - T, m.q);
            }
            __syncthreads();
        }
    } else {
        for (int s = logn - 1; s >= 0; s--) {
            const u32 t = N >> (s + 1), mm = 1u << s;
            for (u32 b = threadIdx.x; b < N / 2; b += blockDim.x) {
                const u32 i = b / t, j = b % t, x = 2 * i * t + j;
                const ulonglong2 W = tw[mm + i];
                const u64 U = a[x], V = a[x + t];
                a[x] = canon2(U + V, m.q);
                a[x + t] = canon2(shoup_mul(U + m.q - V, W.x, W.y, nq), m.q);
            }
            __syncthreads();
        }
        for (u32 i = threadIdx.x; i < N; i += blockDim.x) a[i] = canon2(shoup_mul(a[i], m.ninv, m.ninv_s, nq), m.q);
        __syncthreads();
    }
    AbftOut aout;
    AbftIn acf;
    for (u32 i = threadIdx.x; i < N; i += blockDim.x) {
        base[i] = a[i];
        if (PROT && !INVERSE) aout.add(p.omega[(size_t)mi * N + i], a[i]);
        if (PROT && INVERSE) acf.add(abft_w(i, (logn + 1) / 2), a[i]);        // coefficient side: sum_i w_i a_i
    }
    if (PROT && !INVERSE) { u64 s0, s1; aout.value(s0, s1); abft_commit<5>(p.abft, p.limb0 + limb, 1, s0, s1); }
    if (PROT && INVERSE) { u64 s0, s1; acf.value(s0, s1); abft_commit<4>(p.abft, p.limb0 + limb, 1, s0, s1); }
}

// one thread per (limb, side): recombine the digit sums (Horner, Barrett per step) and clear the counters;
// the even lane of each pair compares the two sides and publishes verdict and sums
__global__ void k_abft_finalize(u64* counters, const B200Mod* mods, u32 L, u32 start_mod, u32 n_limbs, u32* verdict, u64* sums) {
    const u32 t = blockIdx.x * blockDim.x + threadIdx.x;
    const u32 i = t >> 1, side = t & 1;
    u64 r = 0;
    if (i < n_limbs) {
        const B200Mod m = mods[start_mod + i % L];
        u64* c = counters + ((size_t)i * 2 + side) * kAbftStride;
        for (int d = kAbftStride - 1; d >= 0; d--) {   // r < q < 2^60: (r << 26) + c[d] < q 2^64
            const u64 lo = (r << kAbftDigit) + c[d];
            const u64 hi = (r >> (64 - kAbftDigit)) + (lo < c[d]);
            r = barrett128(lo, hi, m);
            c[d] = 0;
        }
    }
    const u64 other = __shfl_xor_sync(0xffffffffu, r, 1);
    if (i < n_limbs && side == 0) {
        if (verdict) verdict[i] = r != other;
        if (sums) { sums[2 * i] = r; sums[2 * i + 1] = other; }
    }
}

// ---------------------------------------------------------------- dispatch
// cudaFuncAttributeMaxDynamicSharedMemorySize is set once per kernel instantiation (not on every launch), and checked
// (the kernel is a non-type template argument: one static per instantiation, not per function-pointer TYPE)
// and per device: the attribute belongs to the device's context)
// set_smem_once<Kernel>(smem): ctx.hpp
// row passes: pairs of polynomials share a CTA (and its twiddle tile); an odd last polynomial runs the one-poly form
// Measured on B200 (profiles/r2e_ablations.txt): the paired form is NOT faster (fwd 107 vs 99 us, inv 103 vs 98.5 us per 288-limb
// batch) -- the row passes are bound by their own instruction stream, not by twiddle traffic or occupancy -- so it stays an
// option (B200FHE_ROW_NP=2) and the one-polynomial form is the default.
static const bool g_row_pairs = getenv("B200FHE_ROW_NP") && atoi(getenv("B200FHE_ROW_NP")) == 2;
template <int BIG, bool PROT, bool INJ>
static int launch_fwd_row(PassArgs a, u32 np, cudaStream_t s) {
    const u32 tiles = (a.N >> kRowLog) / kRowsPerCta, pairs = g_row_pairs ? np / 2 : 0;
    if (pairs) {
        if (set_smem_once<k_fwd_row<BIG, PROT, INJ, 2>>(sizeof(RowSmem<2>)) != cudaSuccess) return -4;
        a.n_poly = pairs * 2; a.poly0 = 0;
        k_fwd_row<BIG, PROT, INJ, 2><<<dim3(tiles * pairs, a.L), kRowSlice * 2, sizeof(RowSmem<2>), s>>>(a);
    }
    if (np - 2 * pairs) {
        if (set_smem_once<k_fwd_row<BIG, PROT, INJ, 1>>(sizeof(RowSmem<1>)) != cudaSuccess) return -4;
        a.n_poly = np - 2 * pairs; a.poly0 = 2 * pairs;
        k_fwd_row<BIG, PROT, INJ, 1><<<dim3(tiles * a.n_poly, a.L), kRowSlice, sizeof(RowSmem<1>), s>>>(a);
    }
    return 0;
}
template <int BIG, bool PROT = false>
static int launch_inv_row(PassArgs a, u32 np, cudaStream_t s) {
    const u32 tiles = (a.N >> kRowLog) / kRowsPerCta, pairs = g_row_pairs ? np / 2 : 0;
    if (a.src) {
        if constexpr (!PROT) {
            if (set_smem_once<k_inv_row<BIG, 1, false, true>>(sizeof(RowSmem<1>)) != cudaSuccess) return -4;
            a.n_poly = np; a.poly0 = 0;
            k_inv_row<BIG, 1, false, true><<<dim3(tiles * np, a.L), kRowSlice, sizeof(RowSmem<1>), s>>>(a);
            return 0;
        }
        return -5;
    }
    if (pairs) {
        if (set_smem_once<k_inv_row<BIG, 2, PROT>>(sizeof(RowSmem<2>)) != cudaSuccess) return -4;
        a.n_poly = pairs * 2; a.poly0 = 0;
        k_inv_row<BIG, 2, PROT><<<dim3(tiles * pairs, a.L), kRowSlice * 2, sizeof(RowSmem<2>), s>>>(a);
    }
    if (np - 2 * pairs) {
        if (set_smem_once<k_inv_row<BIG, 1, PROT>>(sizeof(RowSmem<1>)) != cudaSuccess) return -4;
        a.n_poly = np - 2 * pairs; a.poly0 = 2 * pairs;
        k_inv_row<BIG, 1, PROT><<<dim3(tiles * a.n_poly, a.L), kRowSlice, sizeof(RowSmem<1>), s>>>(a);
    }
    return 0;
}
template <int LC, int LA, int BIG, bool PROT = false, bool INJ = false>
static int launch_fwd2(const PassArgs& a_in, u32 n_limbs, int mask, cudaStream_t s) {
    constexpr int LB = LC - LA;
    const u32 np = n_limbs / a_in.L;
    PassArgs a = a_in;
    a.n_poly = np;
    dim3 gc((kRowLen / kTile) * np, a.L);
    constexpr size_t smem = (LB > 0 ? sizeof(u64) * (kTile << LC) : 0) + sizeof(ulonglong2) * (1 << LC);
    if constexpr (!PROT && !INJ) {
        if (a.map) {                        // scatter list: one polynomial of a.L entries, plain transform only
            if (np != 1) return -5;
            if (mask & 1) {
                if (set_smem_once<k_fwd_col<LC, LA, BIG, false, false, true>>(smem) != cudaSuccess) return -4;
                k_fwd_col<LC, LA, BIG, false, false, true><<<gc, kTile * col_groups(LC, LA), smem, s>>>(a);
            }
            if (mask & 2) {
                if (set_smem_once<k_fwd_row<BIG, false, false, 1, true>>(sizeof(RowSmem<1>)) != cudaSuccess) return -4;
                a.n_poly = 1; a.poly0 = 0;
                k_fwd_row<BIG, false, false, 1, true><<<dim3((a.N >> kRowLog) / kRowsPerCta, a.L), kRowSlice, sizeof(RowSmem<1>), s>>>(a);
            }
            return 0;
        }
    }
    if (a.map) return -5;
    if (mask & 1) {
        if (set_smem_once<k_fwd_col<LC, LA, BIG, PROT, INJ>>(smem) != cudaSuccess) return -4;
        k_fwd_col<LC, LA, BIG, PROT, INJ><<<gc, kTile * col_groups(LC, LA), smem, s>>>(a);
    }
    if (mask & 2) return launch_fwd_row<BIG, PROT, INJ>(a, np, s);
    return 0;
}
template <int LC, int LA, int BIG, bool PROT = false>
static int launch_inv2(const PassArgs& a_in, u32 n_limbs, int mask, cudaStream_t s) {
    constexpr int LB = LC - LA;
    const u32 np = n_limbs / a_in.L;
    if (a_in.map) return -5;              // scatter lists: forward transform only
    PassArgs a = a_in;
    a.n_poly = np;
    dim3 gc((kRowLen / kTile) * np, a.L);
    if (mask & 1) {
        int rc = launch_inv_row<BIG, PROT>(a, np, s);
        if (rc) return rc;
    }
    constexpr size_t smem = (LB > 0 ? sizeof(u64) * (kTile << LC) : 0) + sizeof(ulonglong2) * (1 << LC);
    if (mask & 2) {
        if (set_smem_once<k_inv_col<LC, LA, BIG, PROT>>(smem) != cudaSuccess) return -4;
        k_inv_col<LC, LA, BIG, PROT><<<gc, kTile * col_groups(LC, LA), smem, s>>>(a);
    }
    return 0;
}

template <int BIG>
static int launch_ntt_t(bool inverse, const PassArgs& a, int logn, u32 n_limbs, int mask, cudaStream_t s) {
#define B200_CASE(LOGN, LC, LA)                                              \
    case LOGN:                                                               \
        return inverse ? launch_inv2<LC, LA, BIG>(a, n_limbs, mask, s) : launch_fwd2<LC, LA, BIG>(a, n_limbs, mask, s);
    switch (logn) {
#ifndef B200_FAST_BUILD      // experiment builds instantiate the N = 2^16 kernels only
        B200_CASE(12, 4, 4)
        B200_CASE(13, 5, 3)
        B200_CASE(14, 6, 3)
        B200_CASE(15, 7, 4)
        B200_CASE(17, 9, 5)
#endif
        B200_CASE(16, 8, 4)
    }
#undef B200_CASE
    return -1;
}

// Launch the transform on n_limbs limbs ([n_poly][L][N], limb l of each poly
// uses modulus start_mod + l).  Returns 0, or <0 if logn is unsupported.
int launch_ntt(bool inverse, int big, const PassArgs& a, int logn, u32 n_limbs, int mask, cudaStream_t s) {
    if (n_limbs == 0) return 0;
    if (logn >= 1 && logn <= 11) {
        size_t smem = sizeof(u64) << logn;
        if (inverse) k_small<true, false><<<n_limbs, 256, smem, s>>>(a, logn);
        else k_small<false, false><<<n_limbs, 256, smem, s>>>(a, logn);
        return 0;
    }
    // n_limbs = n_poly * L whole polynomials (limb l of each uses modulus start_mod + l); grid.y = L, grid.x = tiles * n_poly
    // is far below its 2^31 - 1 limit, so the only split is a safety bound of 32768 limbs per launch, cut at polynomial
    // boundaries so the limb -> modulus mapping is kept
    if (a.L == 0 || n_limbs % a.L) return -2;
    const u32 kMaxLimbs = (32768 / a.L ? 32768 / a.L : 1) * a.L;
    for (u32 off = 0; off < n_limbs; off += kMaxLimbs) {
        PassArgs b = a;
        b.data = a.data + (size_t)off * a.N;
        u32 cnt = n_limbs - off < kMaxLimbs ? n_limbs - off : kMaxLimbs;
#ifdef B200_FAST_BUILD
        int rc = big == 2 ? launch_ntt_t<2>(inverse, b, logn, cnt, mask, s) : big ? launch_ntt_t<1>(inverse, b, logn, cnt, mask, s) : -3;
#else
        int rc = big == 2 ? launch_ntt_t<2>(inverse, b, logn, cnt, mask, s)
                 : big  ? launch_ntt_t<1>(inverse, b, logn, cnt, mask, s) : launch_ntt_t<0>(inverse, b, logn, cnt, mask, s);
#endif
        if (rc) return rc;
    }
    return 0;
}

template <int BIG, bool INJ>
static int launch_prot_t(const PassArgs& a, int logn, u32 n_limbs, int mask, cudaStream_t s) {
    switch (logn) {
#ifndef B200_FAST_BUILD
        case 12: return launch_fwd2<4, 4, BIG, true, INJ>(a, n_limbs, mask, s);
        case 13: return launch_fwd2<5, 3, BIG, true, INJ>(a, n_limbs, mask, s);
        case 14: return launch_fwd2<6, 3, BIG, true, INJ>(a, n_limbs, mask, s);
        case 15: return launch_fwd2<7, 4, BIG, true, INJ>(a, n_limbs, mask, s);
        case 17: return launch_fwd2<9, 5, BIG, true, INJ>(a, n_limbs, mask, s);
#endif
        case 16: return launch_fwd2<8, 4, BIG, true, INJ>(a, n_limbs, mask, s);
    }
    return -1;
}

// mask: 3 = both passes, 1 / 2 = first / second pass only (N >= 2^12: the caller corrupts the stored intermediate in between)
int launch_ntt_protected(int big, const PassArgs& a, int logn, u32 n_limbs, cudaStream_t s, int mask) {
    if (n_limbs == 0) return 0;
    if (n_limbs > 32768 || a.L == 0 || n_limbs % a.L) return -2;
    if (logn >= 1 && logn <= 11) {
        if (mask != 3) return -5;
        k_small<false, true><<<n_limbs, 256, sizeof(u64) << logn, s>>>(a, logn);
        return 0;
    }
#ifdef B200_FAST_BUILD
    if (!big || a.n_inj) return -3;
    return big == 2 ? launch_prot_t<2, false>(a, logn, n_limbs, mask, s) : launch_prot_t<1, false>(a, logn, n_limbs, mask, s);
#else
    // the injection variants exist for the general modes only: NEAR moduli are also valid BIG moduli
    if (a.n_inj) return big ? launch_prot_t<1, true>(a, logn, n_limbs, mask, s) : launch_prot_t<0, true>(a, logn, n_limbs, mask, s);
    return big == 2 ? launch_prot_t<2, false>(a, logn, n_limbs, mask, s)
           : big   ? launch_prot_t<1, false>(a, logn, n_limbs, mask, s) : launch_prot_t<0, false>(a, logn, n_limbs, mask, s);
#endif
}

template <int BIG>
static int launch_prot_inv_t(const PassArgs& a, int logn, u32 n_limbs, cudaStream_t s) {
    switch (logn) {
#ifndef B200_FAST_BUILD
        case 12: return launch_inv2<4, 4, BIG, true>(a, n_limbs, 3, s);
        case 13: return launch_inv2<5, 3, BIG, true>(a, n_limbs, 3, s);
        case 14: return launch_inv2<6, 3, BIG, true>(a, n_limbs, 3, s);
        case 15: return launch_inv2<7, 4, BIG, true>(a, n_limbs, 3, s);
        case 17: return launch_inv2<9, 5, BIG, true>(a, n_limbs, 3, s);
#endif
        case 16: return launch_inv2<8, 4, BIG, true>(a, n_limbs, 3, s);
    }
    return -1;
}

// protected inverse transform (a.omega / a.abft set, a.tw / a.rtw the INVERSE tables); counters must be zero on entry
int launch_ntt_protected_inverse(int big, const PassArgs& a, int logn, u32 n_limbs, cudaStream_t s) {
    if (n_limbs == 0) return 0;
    if (n_limbs > 32768 || a.L == 0 || n_limbs % a.L) return -2;
    if (logn >= 1 && logn <= 11) {
        k_small<true, true><<<n_limbs, 256, sizeof(u64) << logn, s>>>(a, logn);
        return 0;
    }
#ifdef B200_FAST_BUILD
    if (!big) return -3;
    return big == 2 ? launch_prot_inv_t<2>(a, logn, n_limbs, s) : launch_prot_inv_t<1>(a, logn, n_limbs, s);
#else
    return big == 2 ? launch_prot_inv_t<2>(a, logn, n_limbs, s)
           : big   ? launch_prot_inv_t<1>(a, logn, n_limbs, s) : launch_prot_inv_t<0>(a, logn, n_limbs, s);
#endif
}

void launch_abft_finalize(u64* counters, const B200Mod* mods, u32 L, u32 start_mod, u32 n_limbs, u32* verdict, u64* sums, cudaStream_t s) {
    if (n_limbs) k_abft_finalize<<<(2 * n_limbs + 127) / 128, 128, 0, s>>>(counters, mods, L, start_mod, n_limbs, verdict, sums);
}

}  // namespace b200
