// Intra-multiplier fold checks of the Barrett multiplier on the device.
//
// The reference studies two low-cost self-checks of a modular multiplier, both as CPU Monte-Carlo simulators:
//   * rfhe_framewk/src/barrett_final.cpp:120-141 (barrett_reduce_signatures) + :175-266 (simulate_once_fold):
//     the low-K-bit SIGNATURES of Barrett's two products mul1 = t mu, mul2 = s q, XOR-ed element with element over a
//     fold group ("instra": lowK(mul1) ^ lowK(mul2); "sum": lowK(c)) and compared between two executions;
//   * rfhe_framewk/src/barrett_test/outer_check.cpp:35-50 (fold_mod_2k_plus_1_u64) + :222-252: the residue of a word
//     modulo 2^k + 1 (alternating sum of its k-bit segments -- shifts and adds, no multiplier), checked per element
//     ("intra") and summed over the vector ("inter").
// Device form: one kernel computes c = Barrett(a b) element-wise exactly as barrett_reduce_signatures does (t = the low
// 64 bits of a b, K = bit_length(q - 1), mu = floor(2^2K / q), one conditional subtraction), and next to it
//   (1) the per-group XOR signatures, so that two executions (or two devices) can be compared at 16 bytes per group
//       instead of 8 bytes per element;
//   (2) the residue code of BOTH products: R(mul1) == R(t) R(mu) and R(mul2) == R(s) R(q) (mod 2^k + 1), where R of the
//       128-bit product is folded on the ALU pipe (SHF / LOP3 / IADD3) and the right-hand side is ONE 32-bit multiply of
//       two (k+1)-bit residues -- a different unit and a different operation from the 64 x 64 -> 128 products (IMAD.WIDE
//       chains) it checks, which is the point of a residue code; and R(t) == R(mul2) + R(c') for the subtraction.
//       "intra" = per element (any element of the group trips bit 0 of the group's verdict), "inter" = the predicted
//       and the observed residues summed over the group and compared once (bit 1; cheaper, blind to cancelling errors,
//       exactly the reference's two columns).
// Faults are injected where barrett_final.cpp:6-7 puts them -- at the multiplication outputs, inside the low K bits:
// kind 1 XORs a mask into the product (SBF / DBF), kind 2 replaces its low K bits (MOF, apply_mof_to_signatures :169-172).
// With propagate = 0 the fault touches only what the checks see (the reference's "signature domain" model, :9); with
// propagate = 1 the corrupted product is the one the arithmetic goes on with (a real multiplier fault).
#include "../../include/b200fhe.h"
#include "ctx.hpp"
#include <vector>
#include <cstdlib>
#include <algorithm>

int b200_set_err(int code, const char* fmt, ...);
void b200_count_launches(unsigned n);

namespace b200 {

struct FoldFault { u64 index; u64 payload; u32 product, kind; };

struct FoldK { u32 k, c, M; };
__host__ __device__ inline FoldK make_foldk(u32 k) {
    u32 c = 2 * k;
    while (2 * c <= 60) c *= 2;
    return FoldK{k, c, (1u << k) + 1u};
}
struct FoldArgs {
    const u64 *a, *b;
    u64* c;
    u64 *sig_intra, *sig_sum;   // [groups] XOR signatures (zeroed before the launch)
    u32* verdict;               // [groups] bit 0 intra residue check, bit 1 inter (written by k_fold_finish), bit 2 subtractor
    unsigned long long* inter;  // [groups][2] predicted / observed residue sums (zeroed)
    const FoldFault* faults;
    const u32* fault_of;        // [n] 1 + index of the first fault on this element, 0 = none (NULL when no faults)
    u64 n, vec_len, groups_per_vec;
    u64 q, mu;
    u32 K, fold, n_faults, propagate;
    FoldK F;
    u32 rmu, rq;                // residues of the two constant operands
};

// residue of the 128-bit word (hi, lo) modulo M = 2^k + 1 (fold_mod_2k_plus_1_u64, outer_check.cpp:35-50, extended to two
// words).  Shifts, masks and adds only (ALU pipe), no loop over the 128 / k segments: 2^(2k) = 1 (mod M), so first add up
// the word's chunks of c = 2k 2^j bits (32 <= c <= 60, at most four of them), halve the chunk width down to 2k bits, and
// finish with the one alternating step low k bits - high k bits.
__device__ __forceinline__ u32 fold128(u64 lo, u64 hi, const FoldK& F) {
    const u32 c = F.c;
    const u64 mc = ((u64)1 << c) - 1;
    u64 s = (lo & mc) + (((lo >> c) | (hi << (64 - c))) & mc) + ((hi >> (2 * c - 64)) & mc);
    if (3 * c < 128) s += hi >> (3 * c - 64);
    for (u32 w = c >> 1; w >= 2 * F.k; w >>= 1) s = (s & (((u64)1 << w) - 1)) + (s >> w);
    const u32 w2 = 2 * F.k;
    const u64 m2 = ((u64)1 << w2) - 1;
    s = (s & m2) + (s >> w2);                      // the halvings leave at most a few carries above 2k bits
    s = (s & m2) + (s >> w2);
    int r = (int)(s & ((1u << F.k) - 1)) - (int)(s >> F.k);
    if (r < 0) r += (int)F.M;
    return (u32)r;
}
__device__ __forceinline__ u32 mulres(u32 x, u32 y, const FoldK& F) {   // x y mod 2^k + 1 for residues x, y <= 2^k
    return fold128((u64)x * y, 0, F);                                 // the check's own small multiply: 32 x 32 -> 64
}

// ALIGNED: fold is a power of two <= 256 dividing vec_len, so a group never straddles a CTA of 256 threads: its
// signature / residue sums are reduced in the warp (and across warps through shared memory) and written with plain
// stores -- no atomics, no clearing pass, the inter comparison done in place.  Otherwise: atomics + k_fold_finish.
// E: elements per thread.  E = 4 (ALIGNED only, 4 | fold, 4 | n, 16-byte aligned operands): a thread's four consecutive elements
// belong to one group, so their signatures are combined in registers first -- a quarter of the cross-lane shuffles (the shuffle
// unit, one warp instruction per cycle per SM, was what bounded the one-element form) and 128-bit loads / stores.
template <bool CHECKS, bool ALIGNED, int E>
__global__ void __launch_bounds__(256) k_barrett_fold(FoldArgs A) {
    static_assert(E == 1 || (E == 4 && ALIGNED), "four elements per thread need the aligned shape");
    const u64 i0 = ((u64)blockIdx.x * blockDim.x + threadIdx.x) * E;
    const bool live = i0 < A.n;
    const u64 maskK = A.K >= 64 ? ~0ull : (((u64)1 << A.K) - 1);
    const FoldK F = A.F;
    u64 sig1 = 0, sig2 = 0;
    u32 bad = 0, pred = 0, obs = 0;
    u64 av[E], bv[E], cv[E];
    if (live) {
        if (E == 4) {
            const ulonglong2 a0 = reinterpret_cast<const ulonglong2*>(A.a + i0)[0], a1 = reinterpret_cast<const ulonglong2*>(A.a + i0)[1];
            const ulonglong2 b0 = reinterpret_cast<const ulonglong2*>(A.b + i0)[0], b1 = reinterpret_cast<const ulonglong2*>(A.b + i0)[1];
            av[0] = a0.x; av[1 % E] = a0.y; av[2 % E] = a1.x; av[3 % E] = a1.y;
            bv[0] = b0.x; bv[1 % E] = b0.y; bv[2 % E] = b1.x; bv[3 % E] = b1.y;
        } else {
            av[0] = A.a[i0]; bv[0] = A.b[i0];
        }
    }
#pragma unroll
    for (int e = 0; e < E; e++) {
      const u64 i = i0 + e;
      if (live) {
        const u64 t = av[e] * bv[e];                                // barrett_final.cpp:239: u64 ti = a * b
        u64 m1lo = t * A.mu, m1hi = __umul64hi(t, A.mu);            // mul1 = ti * mu (:127)
        u64 c1lo = m1lo;                                            // what the checks see
        const u64 c1hi = m1hi;
        const u32 fi = A.fault_of ? A.fault_of[i] : 0;
        for (u32 f = fi; f && f <= A.n_faults && A.faults[f - 1].index == i; f++) {
            const FoldFault X = A.faults[f - 1];
            if (X.product != 1) continue;
            c1lo = X.kind == 1 ? c1lo ^ (X.payload & maskK) : ((c1lo & ~maskK) | (X.payload & maskK));
        }
        if (A.propagate) m1lo = c1lo;
        const u32 sh = 2 * A.K;                                     // si = mul1 >> 2K (:131), 2K <= 120
        const u64 s = sh >= 64 ? (m1hi >> (sh - 64)) : ((m1lo >> sh) | (m1hi << (64 - sh)));
        const u64 shi = sh >= 64 ? 0 : (m1hi >> sh);
        u64 m2lo = s * A.q, m2hi = __umul64hi(s, A.q) + shi * A.q;  // mul2 = si * q (:134)
        u64 c2lo = m2lo;
        const u64 c2hi = m2hi;
        for (u32 f = fi; f && f <= A.n_faults && A.faults[f - 1].index == i; f++) {
            const FoldFault X = A.faults[f - 1];
            if (X.product != 2) continue;
            c2lo = X.kind == 1 ? c2lo ^ (X.payload & maskK) : ((c2lo & ~maskK) | (X.payload & maskK));
        }
        if (A.propagate) m2lo = c2lo;
        u64 ci = t - m2lo;                                          // (:138-139)
        const u64 cpre = ci;
        if (ci >= A.q) ci -= A.q;
        cv[e] = ci;
        sig1 ^= (c1lo & maskK) ^ (c2lo & maskK);                    // "instra" (:251-253)
        sig2 ^= ci & maskK;                                         // "sum"    (:254-257)
        if (CHECKS) {
            const u32 M = F.M;
            const u32 rt = fold128(t, 0, F);
            const u32 r1 = fold128(c1lo, c1hi, F), r2 = fold128(c2lo, c2hi, F);
            const u32 p1 = mulres(rt, A.rmu, F);
            const u32 p2 = mulres(fold128(s, shi, F), A.rq, F);
            if (r1 != p1 || r2 != p2) bad |= 1;
            // subtraction: t == mul2 + c' as integers when mul2 <= t (always, fault-free)
            u32 rs = fold128(m2lo, 0, F) + fold128(cpre, 0, F);
            if (rs >= M) rs -= M;
            if (m2hi == 0 && m2lo <= t && rs != rt) bad |= 4;
            pred += p1 + p2;
            obs += r1 + r2;
        }
      }
    }
    const u64 i = i0;
    if (live) {
        if (E == 4) {
            reinterpret_cast<ulonglong2*>(A.c + i0)[0] = make_ulonglong2(cv[0], cv[1 % E]);
            reinterpret_cast<ulonglong2*>(A.c + i0)[1] = make_ulonglong2(cv[2 % E], cv[3 % E]);
        } else {
            A.c[i0] = cv[0];
        }
        if (CHECKS) { pred %= F.M; obs %= F.M; }
    }
    if (ALIGNED) {
        // n is a multiple of fold and a group (fold / E threads) divides the CTA: every group is complete and lives in this CTA
        const u32 fold = A.fold / E, lane = threadIdx.x & 31;        // threads per group
        u32 s1l = (u32)sig1, s1h = (u32)(sig1 >> 32), s2l = (u32)sig2, s2h = (u32)(sig2 >> 32);
        const u32 span = fold < 32 ? fold : 32;
        for (u32 o = span >> 1; o; o >>= 1) {
            s1l ^= __shfl_xor_sync(0xffffffffu, s1l, o); s1h ^= __shfl_xor_sync(0xffffffffu, s1h, o);
            s2l ^= __shfl_xor_sync(0xffffffffu, s2l, o); s2h ^= __shfl_xor_sync(0xffffffffu, s2h, o);
            if (CHECKS) {
                bad |= __shfl_xor_sync(0xffffffffu, bad, o);
                pred += __shfl_xor_sync(0xffffffffu, pred, o);       // < 32 M < 2^32
                obs += __shfl_xor_sync(0xffffffffu, obs, o);
            }
        }
        if (fold <= 32) {
            if (live && (lane & (fold - 1)) == 0) {
                const u64 g = i / A.fold;
                A.sig_intra[g] = ((u64)s1h << 32) | s1l;
                A.sig_sum[g] = ((u64)s2h << 32) | s2l;
                if (CHECKS) A.verdict[g] = bad | (pred % F.M != obs % F.M ? 2u : 0u);
            }
        } else {
            __shared__ u32 sm[8][7];
            const u32 warp = threadIdx.x >> 5;
            if (lane == 0) {
                sm[warp][0] = s1l; sm[warp][1] = s1h; sm[warp][2] = s2l; sm[warp][3] = s2h;
                sm[warp][4] = bad; sm[warp][5] = pred % F.M; sm[warp][6] = obs % F.M;
            }
            __syncthreads();
            const u32 wpg = fold >> 5, ngrp = 256 / fold;            // warps per group, groups per CTA
            if (threadIdx.x < ngrp) {
                const u64 g = ((u64)blockIdx.x * 256) / fold + threadIdx.x;
                if (g * A.fold < A.n) {
                    u32 v[7] = {0, 0, 0, 0, 0, 0, 0};
                    for (u32 w = threadIdx.x * wpg; w < (threadIdx.x + 1) * wpg; w++) {
                        v[0] ^= sm[w][0]; v[1] ^= sm[w][1]; v[2] ^= sm[w][2]; v[3] ^= sm[w][3];
                        v[4] |= sm[w][4]; v[5] += sm[w][5]; v[6] += sm[w][6];
                    }
                    A.sig_intra[g] = ((u64)v[1] << 32) | v[0];
                    A.sig_sum[g] = ((u64)v[3] << 32) | v[2];
                    if (CHECKS) A.verdict[g] = v[4] | (v[5] % F.M != v[6] % F.M ? 2u : 0u);
                }
            }
        }
        return;
    }
    // general shapes: a warp that lies inside one group reduces first, one atomic per warp; otherwise per element
    const u64 vec = live ? i / A.vec_len : 0, within = live ? i % A.vec_len : 0;
    const u64 g = vec * A.groups_per_vec + within / A.fold;
    const u64 g0 = __shfl_sync(0xffffffffu, g, 0);
    const bool uniform = __all_sync(0xffffffffu, live && g == g0);
    if (uniform) {
        const u32 lo = __reduce_xor_sync(0xffffffffu, (u32)sig1), hi = __reduce_xor_sync(0xffffffffu, (u32)(sig1 >> 32));
        const u32 lo2 = __reduce_xor_sync(0xffffffffu, (u32)sig2), hi2 = __reduce_xor_sync(0xffffffffu, (u32)(sig2 >> 32));
        u32 anybad = 0, ps = 0, os = 0;
        if (CHECKS) {
            anybad = __reduce_or_sync(0xffffffffu, bad);
            ps = __reduce_add_sync(0xffffffffu, pred);
            os = __reduce_add_sync(0xffffffffu, obs);
        }
        if ((threadIdx.x & 31) == 0) {
            atomicXor((unsigned long long*)&A.sig_intra[g], ((u64)hi << 32) | lo);
            atomicXor((unsigned long long*)&A.sig_sum[g], ((u64)hi2 << 32) | lo2);
            if (CHECKS) {
                if (anybad) atomicOr(&A.verdict[g], anybad);
                atomicAdd(&A.inter[2 * g], (unsigned long long)(ps % F.M));
                atomicAdd(&A.inter[2 * g + 1], (unsigned long long)(os % F.M));
            }
        }
    } else if (live) {
        atomicXor((unsigned long long*)&A.sig_intra[g], sig1);
        atomicXor((unsigned long long*)&A.sig_sum[g], sig2);
        if (CHECKS) {
            if (bad) atomicOr(&A.verdict[g], bad);
            atomicAdd(&A.inter[2 * g], (unsigned long long)pred);
            atomicAdd(&A.inter[2 * g + 1], (unsigned long long)obs);
        }
    }
}

// inter check of the general path: predicted vs observed residue sums of each group, compared once (outer_check.cpp:236-247)
__global__ void k_fold_finish(u32* verdict, const unsigned long long* inter, u64 groups, u32 k) {
    const u64 g = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= groups) return;
    const u32 M = (1u << k) + 1u;
    if (inter[2 * g] % M != inter[2 * g + 1] % M) verdict[g] |= 2;
}

// plain element-wise Barrett product, the unchecked baseline the overhead is measured against
__global__ void __launch_bounds__(256) k_barrett_plain(const u64* a, const u64* b, u64* c, u64 n, u64 q, u64 mu, u32 K) {
    const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u64 t = a[i] * b[i];
    const u64 m1lo = t * mu, m1hi = __umul64hi(t, mu);
    const u32 sh = 2 * K;
    const u64 s = sh >= 64 ? (m1hi >> (sh - 64)) : ((m1lo >> sh) | (m1hi << (64 - sh)));
    u64 ci = t - s * q;
    if (ci >= q) ci -= q;
    c[i] = ci;
}

}  // namespace b200

using namespace b200;

static int bitlen(uint64_t v) { int b = 0; while (v) { b++; v >>= 1; } return b; }

extern "C" int b200fhe_barrett_params(uint64_t q, int* K, uint64_t* mu) {
    if (q < 3 || (q >> 60)) return b200_set_err(B200FHE_EINVAL, "barrett_params: need 3 <= q < 2^60");
    const int k = bitlen(q - 1);                                     // make_barrett_ctx (barrett_final.cpp:74-80)
    const unsigned __int128 num = (unsigned __int128)1 << (2 * k);
    if (K) *K = k;
    if (mu) *mu = (uint64_t)(num / q);
    return 0;
}

// Device-resident form.  mode bit 0: compute the residue checks (verdicts); 0 = signatures only.  mode = -1: the plain
// unchecked product (baseline for the overhead measurement; signature / verdict pointers ignored).
extern "C" int b200fhe_barrett_fold(uint64_t q, const uint64_t* d_a, const uint64_t* d_b, uint64_t* d_c, size_t n_vec, size_t vec_len,
                                    int fold, int k, int mode, int propagate, const b200fhe_fold_fault* h_faults, size_t n_faults,
                                    uint64_t* d_sig_intra, uint64_t* d_sig_sum, uint32_t* d_verdict, void* stream) {
    int K; uint64_t mu;
    if (int rc = b200fhe_barrett_params(q, &K, &mu)) return rc;
    const size_t n = n_vec * vec_len;
    if (!d_a || !d_b || !d_c || !n) return b200_set_err(B200FHE_EINVAL, "barrett_fold: null or empty operands");
    cudaStream_t s = (cudaStream_t)stream;
    const unsigned blocks = (unsigned)((n + 255) / 256);
    if (mode < 0) {
        k_barrett_plain<<<blocks, 256, 0, s>>>((const u64*)d_a, (const u64*)d_b, (u64*)d_c, n, q, mu, (u32)K);
        b200_count_launches(1);
        cudaError_t e = cudaGetLastError();
        return e == cudaSuccess ? 0 : b200_set_err(B200FHE_ECUDA, "barrett_fold: %s", cudaGetErrorString(e));
    }
    const bool checks = mode & 1;
    if (fold < 1 || (size_t)fold > vec_len) return b200_set_err(B200FHE_EINVAL, "barrett_fold: fold must be in [1, vec_len]");
    if (checks && (k < 2 || k > 26)) return b200_set_err(B200FHE_EINVAL, "barrett_fold: residue width k must be in [2, 26]");
    if (!d_sig_intra || !d_sig_sum || (checks && !d_verdict)) return b200_set_err(B200FHE_EINVAL, "barrett_fold: null outputs");
    const size_t gpv = (vec_len + fold - 1) / fold, groups = gpv * n_vec;
    const bool pow2 = (fold & (fold - 1)) == 0 && vec_len % fold == 0;
    const bool four = pow2 && fold >= 4 && fold <= 1024 && n % 4 == 0 &&
                      ((((uintptr_t)d_a) | ((uintptr_t)d_b) | ((uintptr_t)d_c)) & 15) == 0 && !getenv("B200FHE_FOLD_E1");
    const bool aligned = four || (pow2 && fold <= 256);
    FoldArgs A{};
    A.a = (const u64*)d_a; A.b = (const u64*)d_b; A.c = (u64*)d_c; A.sig_intra = (u64*)d_sig_intra; A.sig_sum = (u64*)d_sig_sum; A.verdict = d_verdict;
    A.n = n; A.vec_len = vec_len; A.groups_per_vec = gpv; A.q = q; A.mu = mu; A.K = (u32)K; A.fold = (u32)fold;
    A.n_faults = (u32)n_faults; A.propagate = propagate ? 1 : 0;
    if (checks) {
        A.F = make_foldk((u32)k);
        A.rmu = (u32)(mu % A.F.M); A.rq = (u32)(q % A.F.M);
    }
    cudaError_t e = cudaSuccess;
    unsigned long long* d_inter = nullptr; FoldFault* d_f = nullptr; u32* d_of = nullptr;
    if (!aligned) {
        e = cudaMemsetAsync(d_sig_intra, 0, groups * 8, s);
        if (e == cudaSuccess) e = cudaMemsetAsync(d_sig_sum, 0, groups * 8, s);
        if (e == cudaSuccess && checks) {
            e = cudaMemsetAsync(d_verdict, 0, groups * 4, s);
            if (e == cudaSuccess) e = cudaMallocAsync(&d_inter, groups * 16, s);
            if (e == cudaSuccess) e = cudaMemsetAsync(d_inter, 0, groups * 16, s);
            A.inter = d_inter;
        }
    }
    std::vector<FoldFault> fl;
    std::vector<u32> of;
    if (e == cudaSuccess && n_faults) {
        if (!h_faults) return b200_set_err(B200FHE_EINVAL, "barrett_fold: null fault list");
        fl.resize(n_faults);
        for (size_t i = 0; i < n_faults; i++) {
            if (h_faults[i].index >= n || h_faults[i].product < 1 || h_faults[i].product > 2 || h_faults[i].kind < 1 || h_faults[i].kind > 2)
                return b200_set_err(B200FHE_EINVAL, "barrett_fold: fault %zu out of range", i);
            fl[i] = FoldFault{h_faults[i].index, h_faults[i].payload, h_faults[i].product, h_faults[i].kind};
        }
        std::stable_sort(fl.begin(), fl.end(), [](const FoldFault& x, const FoldFault& y) { return x.index < y.index; });
        of.assign(n, 0);
        for (size_t i = n_faults; i-- > 0;) of[fl[i].index] = (u32)i + 1;
        e = cudaMallocAsync(&d_f, sizeof(FoldFault) * n_faults, s);
        if (e == cudaSuccess) e = cudaMallocAsync(&d_of, n * 4, s);
        if (e == cudaSuccess) e = cudaMemcpyAsync(d_f, fl.data(), sizeof(FoldFault) * n_faults, cudaMemcpyHostToDevice, s);
        if (e == cudaSuccess) e = cudaMemcpyAsync(d_of, of.data(), n * 4, cudaMemcpyHostToDevice, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);          // fl / of are pageable host vectors
        A.faults = d_f; A.fault_of = d_of;
    }
    if (e == cudaSuccess) {
        const unsigned blocks4 = (unsigned)((n / 4 + 255) / 256);
        if (checks && four) k_barrett_fold<true, true, 4><<<blocks4, 256, 0, s>>>(A);
        else if (four) k_barrett_fold<false, true, 4><<<blocks4, 256, 0, s>>>(A);
        else if (checks && aligned) k_barrett_fold<true, true, 1><<<blocks, 256, 0, s>>>(A);
        else if (checks) k_barrett_fold<true, false, 1><<<blocks, 256, 0, s>>>(A);
        else if (aligned) k_barrett_fold<false, true, 1><<<blocks, 256, 0, s>>>(A);
        else k_barrett_fold<false, false, 1><<<blocks, 256, 0, s>>>(A);
        b200_count_launches(1);
        e = cudaGetLastError();
        if (e == cudaSuccess && checks && !aligned) {
            k_fold_finish<<<(unsigned)((groups + 255) / 256), 256, 0, s>>>(d_verdict, d_inter, groups, (u32)k);
            b200_count_launches(1);
            e = cudaGetLastError();
        }
    }
    if (d_inter) cudaFreeAsync(d_inter, s);
    if (d_f) cudaFreeAsync(d_f, s);
    if (d_of) cudaFreeAsync(d_of, s);
    if (e != cudaSuccess) return b200_set_err(B200FHE_ECUDA, "barrett_fold: %s", cudaGetErrorString(e));
    return 0;
}

extern "C" int b200fhe_barrett_fold_host(uint64_t q, const uint64_t* h_a, const uint64_t* h_b, uint64_t* h_c, size_t n_vec, size_t vec_len,
                                         int fold, int k, int mode, int propagate, const b200fhe_fold_fault* h_faults, size_t n_faults,
                                         uint64_t* h_sig_intra, uint64_t* h_sig_sum, uint32_t* h_verdict, int device) {
    if (!h_a || !h_b || !h_c || fold < 1 || !vec_len) return b200_set_err(B200FHE_EINVAL, "barrett_fold_host: bad args");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return b200_set_err(B200FHE_ECUDA, "barrett_fold_host: %s", cudaGetErrorString(e));
    const size_t n = n_vec * vec_len, groups = ((vec_len + fold - 1) / fold) * n_vec;
    uint64_t *d_a = nullptr, *d_b = nullptr, *d_c = nullptr, *d_s1 = nullptr, *d_s2 = nullptr;
    uint32_t* d_v = nullptr;
    e = cudaMalloc(&d_a, n * 8);
    if (e == cudaSuccess) e = cudaMalloc(&d_b, n * 8);
    if (e == cudaSuccess) e = cudaMalloc(&d_c, n * 8);
    if (e == cudaSuccess) e = cudaMalloc(&d_s1, groups * 8);
    if (e == cudaSuccess) e = cudaMalloc(&d_s2, groups * 8);
    if (e == cudaSuccess) e = cudaMalloc(&d_v, groups * 4);
    if (e == cudaSuccess) e = cudaMemcpy(d_a, h_a, n * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_b, h_b, n * 8, cudaMemcpyHostToDevice);
    int rc = 0;
    if (e == cudaSuccess) {
        rc = b200fhe_barrett_fold(q, d_a, d_b, d_c, n_vec, vec_len, fold, k, mode, propagate, h_faults, n_faults, d_s1, d_s2, d_v, nullptr);
        if (!rc) e = cudaDeviceSynchronize();
    }
    if (!rc && e == cudaSuccess) e = cudaMemcpy(h_c, d_c, n * 8, cudaMemcpyDeviceToHost);
    if (!rc && e == cudaSuccess && mode >= 0 && h_sig_intra) e = cudaMemcpy(h_sig_intra, d_s1, groups * 8, cudaMemcpyDeviceToHost);
    if (!rc && e == cudaSuccess && mode >= 0 && h_sig_sum) e = cudaMemcpy(h_sig_sum, d_s2, groups * 8, cudaMemcpyDeviceToHost);
    if (!rc && e == cudaSuccess && mode > 0 && h_verdict) e = cudaMemcpy(h_verdict, d_v, groups * 4, cudaMemcpyDeviceToHost);
    cudaFree(d_a); cudaFree(d_b); cudaFree(d_c); cudaFree(d_s1); cudaFree(d_s2); cudaFree(d_v);
    if (rc) return rc;
    if (e != cudaSuccess) return b200_set_err(B200FHE_ECUDA, "barrett_fold_host: %s", cudaGetErrorString(e));
    return 0;
}
