// RNS ciphertext x plaintext multiply-accumulate in the NTT domain (config C3):
//   out[p][l][i] = sum_{k < T} ct_k[p][l][i] * pt_k[l][i]  mod q_l
// Experiment definition: lattigo-error-demo/seal_pmult.cpp:52 (multiply_plain),
// accumulate pattern profile_framewk/src/matmul_ckks.cpp:90-107, SEAL kernel
// dyadic_product_coeffmod; GPU call site of the ct x ct flavour
// reliability_test/dotprod_test.cu:113.
//
// Pure streaming kernel, HBM-bound: every ct word is read once, every pt word
// once per call (shared by the P polynomials of a ciphertext inside the thread),
// products are accumulated lazily in 192 bits and reduced ONCE per output
// coefficient (two 128-bit Barrett steps), so arbitrary 64-bit input words
// (bit-flipped residues, dotprod_test.cu:38-61) are handled exactly: the result
// equals sum (ct mod q)(pt mod q) mod q.
#include "ctx.hpp"

namespace b200 {

constexpr int kPmultTerms = 64;   // terms per launch (pointer table in kernel parameters)
struct PmultPtrs {
    const u64* ct[kPmultTerms];   // each [P][L][N]
    const u64* pt[kPmultTerms];   // each [L][N]
};

struct Acc192 {
    u64 a0, a1, a2;
};
__device__ __forceinline__ void acc_mad(Acc192& A, u64 x, u64 y) {
    const u64 lo = x * y, hi = __umul64hi(x, y);
    asm("add.cc.u64 %0, %0, %3;\n\taddc.cc.u64 %1, %1, %4;\n\taddc.u64 %2, %2, 0;"
        : "+l"(A.a0), "+l"(A.a1), "+l"(A.a2)
        : "l"(lo), "l"(hi));
}
// (a2:a1:a0) mod q, any 192-bit value.  Short sums of in-range operands (the key-switch's inner product: dnum terms) stay below
// 2^(2 bits + 4) and take the pseudo-Mersenne fold of reduce128 (5 IMAD.WIDE) instead of two 128-bit Barretts (~34): with 3 terms
// per output the reduction was three quarters of the kernel's multiplier work.
__device__ __forceinline__ u64 acc_reduce(const Acc192& A, const B200Mod& m) {
    if (A.a2 == 0) return reduce128(A.a0, A.a1, m);
    const u64 r1 = barrett128(A.a1, A.a2, m);        // (a2:a1) mod q  < q
    return barrett128(A.a0, r1, m);                  // (r1:a0) < q 2^64
}

// one thread = two consecutive coefficients of one limb, all P polynomials.
// TT > 0: the number of terms at compile time (the key-switch's inner product has dnum = 2..4 of them): every load of the thread is
// issued before the first product, instead of two terms at a time.
template <int P, int TT = 0>
__global__ void __launch_bounds__(256) k_pmult(const __grid_constant__ PmultPtrs ptrs, int T, u64* out, const B200Mod* mods,
                                               u32 N, u32 L, u32 start_mod, int accumulate) {
    const u32 limb = blockIdx.y;
    const B200Mod m = mods[start_mod + limb];
    const size_t i2 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;   // index of the coefficient pair
    if (i2 * 2 >= N) return;
    const size_t off = ((size_t)limb * N) / 2 + i2;                    // in ulonglong2 units inside one [L][N] block
    const size_t poly2 = ((size_t)L * N) / 2;
    Acc192 acc[P][2];
#pragma unroll
    for (int p = 0; p < P; p++) {
        acc[p][0] = Acc192{0, 0, 0};
        acc[p][1] = Acc192{0, 0, 0};
    }
    if (accumulate) {
#pragma unroll
        for (int p = 0; p < P; p++) {
            const ulonglong2 o = reinterpret_cast<const ulonglong2*>(out)[p * poly2 + off];
            acc[p][0].a0 = o.x;
            acc[p][1].a0 = o.y;
        }
    }
    if constexpr (TT > 0) {
        ulonglong2 w[TT], c[TT][P];
#pragma unroll
        for (int k = 0; k < TT; k++) {
            w[k] = __ldcs(reinterpret_cast<const ulonglong2*>(ptrs.pt[k]) + off);
#pragma unroll
            for (int p = 0; p < P; p++) c[k][p] = __ldcs(reinterpret_cast<const ulonglong2*>(ptrs.ct[k]) + p * poly2 + off);
        }
#pragma unroll
        for (int k = 0; k < TT; k++)
#pragma unroll
            for (int p = 0; p < P; p++) {
                acc_mad(acc[p][0], c[k][p].x, w[k].x);
                acc_mad(acc[p][1], c[k][p].y, w[k].y);
            }
    } else {
#pragma unroll 2
        for (int k = 0; k < T; k++) {
            const ulonglong2 w = __ldcs(reinterpret_cast<const ulonglong2*>(ptrs.pt[k]) + off);
            ulonglong2 c[P];
#pragma unroll
            for (int p = 0; p < P; p++) c[p] = __ldcs(reinterpret_cast<const ulonglong2*>(ptrs.ct[k]) + p * poly2 + off);
#pragma unroll
            for (int p = 0; p < P; p++) {
                acc_mad(acc[p][0], c[p].x, w.x);
                acc_mad(acc[p][1], c[p].y, w.y);
            }
        }
    }
#pragma unroll
    for (int p = 0; p < P; p++)
        reinterpret_cast<ulonglong2*>(out)[p * poly2 + off] = make_ulonglong2(acc_reduce(acc[p][0], m), acc_reduce(acc[p][1], m));
}

// The key-switch's inner product over the Q limbs fused with ModDown's last step: out[p][l][i] =
// (sum_k ct_k[p][l][i] pt_k[l][i] - conv[p][l][i]) Pinv_l mod q_l for the two polynomials of a switch.  The accumulated product never
// goes to memory (2 n_q limbs written and read back otherwise) and k_moddown_combine's launch disappears; the arithmetic is the
// same sequence (reduce to [0,q), subtract the canonical conv, one Shoup product), so the output bits are those of the two-kernel form.
// c_own / alpha (optional): term k's plaintext-side operand for the limbs of digit k (limb / alpha == k) is read from c_own [L][N]
// instead of pt[k] -- a digit's own limbs ARE the ciphertext's, so the key-switch need not copy them into the raised digits.
template <int TT>
__global__ void __launch_bounds__(256) k_pmult_moddown(const __grid_constant__ PmultPtrs ptrs, const u64* __restrict__ conv, u64* __restrict__ out,
                                                       const B200Mod* mods, const ulonglong2* __restrict__ pinv, u32 N, u32 L,
                                                       const u64* __restrict__ c_own, u32 alpha) {
    const u32 limb = blockIdx.y;
    const B200Mod m = mods[limb];
    const size_t i2 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i2 * 2 >= N) return;
    const size_t off = ((size_t)limb * N) / 2 + i2, poly2 = ((size_t)L * N) / 2;
    ulonglong2 w[TT], c[TT][2], b[2];
#pragma unroll
    for (int k = 0; k < TT; k++) {
        const u64* wsrc = (c_own && limb / alpha == (u32)k) ? c_own : ptrs.pt[k];
        w[k] = __ldcs(reinterpret_cast<const ulonglong2*>(wsrc) + off);
#pragma unroll
        for (int p = 0; p < 2; p++) c[k][p] = __ldcs(reinterpret_cast<const ulonglong2*>(ptrs.ct[k]) + p * poly2 + off);
    }
#pragma unroll
    for (int p = 0; p < 2; p++) b[p] = __ldcs(reinterpret_cast<const ulonglong2*>(conv) + p * poly2 + off);
    const ulonglong2 pv = pinv[limb];
    const u64 q = m.q, nq = 0 - q;
#pragma unroll
    for (int p = 0; p < 2; p++) {
        Acc192 a0{0, 0, 0}, a1{0, 0, 0};
#pragma unroll
        for (int k = 0; k < TT; k++) {
            acc_mad(a0, c[k][p].x, w[k].x);
            acc_mad(a1, c[k][p].y, w[k].y);
        }
        const u64 rx = acc_reduce(a0, m), ry = acc_reduce(a1, m);
        const u64 dx = rx - b[p].x + (rx < b[p].x ? q : 0), dy = ry - b[p].y + (ry < b[p].y ? q : 0);
        ulonglong2 r;
        r.x = canon2(shoup_mul(dx, pv.x, pv.y, nq), q);
        r.y = canon2(shoup_mul(dy, pv.x, pv.y, nq), q);
        reinterpret_cast<ulonglong2*>(out)[p * poly2 + off] = r;
    }
}

// ct[k] [2][L][N], pt[k] [L][N], conv [2][L][N], out [2][L][N], moduli 0..L-1; T in 1..4 (else -1: the caller keeps the two-kernel form)
int launch_pmult_moddown(const u64* const* ct, const u64* const* pt, size_t T, const u64* conv, u64* out, const B200Mod* mods,
                         const ulonglong2* pinv, u32 N, u32 L, cudaStream_t s, const u64* c_own, u32 alpha) {
    if (c_own && !alpha) return -1;
    if (T < 1 || T > 4 || N < 2) return -1;
    PmultPtrs ptrs;
    for (int k = 0; k < kPmultTerms; k++) {
        ptrs.ct[k] = (size_t)k < T ? ct[k] : nullptr;
        ptrs.pt[k] = (size_t)k < T ? pt[k] : nullptr;
    }
    dim3 grid((N / 2 + 255) / 256, L);
    switch (T) {
        case 1: k_pmult_moddown<1><<<grid, 256, 0, s>>>(ptrs, conv, out, mods, pinv, N, L, c_own, alpha); break;
        case 2: k_pmult_moddown<2><<<grid, 256, 0, s>>>(ptrs, conv, out, mods, pinv, N, L, c_own, alpha); break;
        case 3: k_pmult_moddown<3><<<grid, 256, 0, s>>>(ptrs, conv, out, mods, pinv, N, L, c_own, alpha); break;
        default: k_pmult_moddown<4><<<grid, 256, 0, s>>>(ptrs, conv, out, mods, pinv, N, L, c_own, alpha); break;
    }
    return 0;
}

// ct[k]: T device pointers to [P][L][N]; pt[k]: T device pointers to [L][N]; out [P][L][N].
int launch_pmult(const u64* const* ct, const u64* const* pt, size_t T, u64* out, const B200Mod* mods, u32 N, u32 P, u32 L,
                 u32 start_mod, int accumulate, cudaStream_t s, unsigned long long* launches) {
    if (N < 2) return -1;
    const u32 threads = 256, pairs = N / 2;
    dim3 grid((pairs + threads - 1) / threads, L);
    for (u32 p0 = 0; p0 < P;) {
        const u32 pc = (P - p0 >= 2) ? 2 : 1;
        const size_t poff = (size_t)p0 * L * N;
        for (size_t k0 = 0; k0 < T || (T == 0 && k0 == 0); k0 += kPmultTerms) {
            PmultPtrs ptrs;
            const int tc = (int)((T - k0) < (size_t)kPmultTerms ? (T - k0) : kPmultTerms);
            for (int k = 0; k < kPmultTerms; k++) {
                ptrs.ct[k] = k < tc ? ct[k0 + k] + poff : nullptr;
                ptrs.pt[k] = k < tc ? pt[k0 + k] : nullptr;
            }
            const int acc_in = accumulate || k0 != 0;
#define B200_PMULT_LAUNCH(PP, TTT) k_pmult<PP, TTT><<<grid, threads, 0, s>>>(ptrs, tc, out + poff, mods, N, L, start_mod, acc_in)
            if (pc == 2) {
                switch (tc) {
                    case 2: B200_PMULT_LAUNCH(2, 2); break;
                    case 3: B200_PMULT_LAUNCH(2, 3); break;
                    case 4: B200_PMULT_LAUNCH(2, 4); break;
                    default: B200_PMULT_LAUNCH(2, 0);
                }
            } else {
                switch (tc) {
                    case 2: B200_PMULT_LAUNCH(1, 2); break;
                    case 3: B200_PMULT_LAUNCH(1, 3); break;
                    case 4: B200_PMULT_LAUNCH(1, 4); break;
                    default: B200_PMULT_LAUNCH(1, 0);
                }
            }
#undef B200_PMULT_LAUNCH
            if (launches) ++*launches;
            if (T == 0) break;
        }
        p0 += pc;
    }
    return 0;
}

}  // namespace b200
