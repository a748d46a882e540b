// The protected four-step NTT the reference EVALUATED (rfhe_framewk/src/ntt_test/relia_ntt_sim.cpp:295-361
// four_step_ntt), with its four detectors, on the device -- one CTA per trial:
//   stage 1  n1 column transforms of length n2 (:310-316, radix-2 DIT after a bit-reversal, :169-206)
//            + batch check   sum_i w^_i s_in_i == sum_i w_i s_out_i  on the column sums (:235-274, check_cols)
//   stage 2  element-wise twiddle multiply by wN^(r c) (:319-335, :208-232)
//            + intra check   fold checksum of every product (:135-142, :222-227)
//            + inter check   sum phi_r post == sum (phi_r T) pre (:276-292)
//   stage 3  transpose and n2 row transforms of length n1 (:337-352) + batch check on the row sums (:354)
//   verdict  detected = !b1 || intra || !inter || !b2 (:359)
// Faults corrupt the product v = w * A[k + half] of one butterfly (stages 1, 3) or one twiddle product (stage 2),
// addressed by the simulator's own running operation index (:185-189, :216-220): op in [0, s1) stage 1,
// [s1, s1 + N) stage 2, then stage 3.  kind 1: v = (v XOR payload) mod q (SBF / DBF, :114-127), kind 2: v = payload mod q
// (MLF, :128-130).  The random check weights (w of batch 1, phi, w of batch 2: rng.rand_q() in the simulator) are inputs,
// so a trial is a deterministic function of its arguments and can be compared bit for bit with the CPU restatement.
#include "../../include/b200fhe.h"
#include "ctx.hpp"
#include "numth.hpp"
#include <vector>
#include <algorithm>

int b200_set_err(int code, const char* fmt, ...);
B200Mod b200_make_mod(uint64_t q, uint64_t itw1, uint64_t n_inv);
void b200_count_launches(unsigned n);

namespace b200 {

struct ReliaFault { u64 op; u64 payload; u32 kind, pad; };

__device__ __forceinline__ u64 mm(u64 a, u64 b, const B200Mod& m) { return mulmod_barrett(a, b, m); }
__device__ __forceinline__ u64 addm(u64 a, u64 b, u64 q) { u64 s = a + b; return s >= q ? s - q : s; }
__device__ __forceinline__ u32 brev(u32 x, int bits) { return bits ? __brev(x) >> (32 - bits) : 0; }
__device__ __forceinline__ u64 fold_sum(u64 v, int w, u64 mask) {      // calc_fold_checksum (:135-142)
    u64 s = 0;
    while (v) { s += v & mask; v >>= w; }
    return s & mask;
}
__device__ __forceinline__ u64 corrupt(u64 v, long long op, const ReliaFault* f, u32 nf, const B200Mod& m) {
    for (u32 i = 0; i < nf; i++)
        if ((long long)f[i].op == op) v = f[i].kind == 1 ? (v ^ f[i].payload) % m.q : f[i].payload % m.q;
    return v;
}

// n_sub transforms of length len = 2^lg on contiguous blocks of x (in place), natural order in and out:
// bit-reversal permutation, then radix-2 DIT stages with w = root^(k len/length) from the power table pw (stride ps).
// op numbering: base + sub * (len/2) lg + stage * (len/2) + group * half + k, as the simulator's loops run.
__device__ void sub_ntts(u64* x, u64* tmp, u32 n_sub, int lg, const u64* pw, u32 ps, long long base, const ReliaFault* f, u32 nf,
                         const B200Mod& m) {
    const u32 len = 1u << lg, tot = n_sub << lg;
    for (u32 i = threadIdx.x; i < tot; i += blockDim.x) tmp[(i & ~(len - 1)) | brev(i & (len - 1), lg)] = x[i];
    __syncthreads();
    for (u32 i = threadIdx.x; i < tot; i += blockDim.x) x[i] = tmp[i];
    __syncthreads();
    const u32 hb = len >> 1;
    for (int s = 0; s < lg; s++) {
        const u32 half = 1u << s;
        for (u32 b = threadIdx.x; b < n_sub * hb; b += blockDim.x) {
            const u32 sub = b / hb, within = b % hb, grp = within >> s, k = within & (half - 1);
            u64* p = x + ((size_t)sub << lg) + (grp << (s + 1)) + k;
            const u64 w = pw[(size_t)(k << (lg - 1 - s)) * ps];
            const u64 u = p[0];
            u64 v = mm(p[half], w, m);
            if (nf) v = corrupt(v, base + (long long)sub * hb * lg + (long long)s * hb + within, f, nf, m);
            p[0] = addm(u, v, m.q);
            p[half] = addm(u, m.q - v, m.q);
        }
        __syncthreads();
    }
}

__device__ u64 block_sum(u64 v, u64 q, u64* red) {
    for (int o = 16; o; o >>= 1) { v = addm(v, __shfl_down_sync(0xffffffffu, v, o), q); }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    u64 r = 0;
    for (u32 i = 0; i < (blockDim.x >> 5); i++) r = addm(r, red[i], q);
    return r;
}

// sm: pw[N] | a[N] | b[N] | c[N] | s_in[max(n1,n2)] | s_out[..] | wv[..] | wh[..] | red[32]
__global__ void __launch_bounds__(256) k_relia_four_step(const u64* __restrict__ in, u64* __restrict__ out, const u64* __restrict__ wts, u32 N, int lg1,
                                                         int lg2, B200Mod m, const u64* __restrict__ pw_g, int fold_bits, const ReliaFault* faults,
                                                         const u32* fault_begin, u32* flags) {
    extern __shared__ u64 sm[];
    const u32 n1 = 1u << lg1, n2 = 1u << lg2, nm = n1 > n2 ? n1 : n2;
    u64 *pw = sm, *a = sm + N, *b = a + N, *c = b + N, *s_in = c + N, *s_out = s_in + nm, *wv = s_out + nm, *wh = wv + nm, *red = wh + nm;
    const u32 tr = blockIdx.x;
    const u64 q = m.q;
    const ReliaFault* f = faults + fault_begin[tr];
    const u32 nf = fault_begin[tr + 1] - fault_begin[tr];
    const u64* w_b1 = wts + (size_t)tr * (2 * n2 + n1), *phi = w_b1 + n2, *w_b2 = phi + n2;
    for (u32 i = threadIdx.x; i < N; i += blockDim.x) {
        pw[i] = pw_g[i];
        const u64 v = in[(size_t)tr * N + i] % q;
        a[i] = v; b[i] = v;
    }
    __syncthreads();
    u32 flag = 0;
    const long long hb1 = (long long)(n2 >> 1) * lg2, s1 = (long long)n1 * hb1, s3 = s1 + N;
    // ---- stage 1: n1 blocks of n2 (root w_n2 = wN^n1) ----
    sub_ntts(b, c, n1, lg2, pw, n1, 0, f, nf, m);
    {   // batch check, check_cols: s[r] = sum_c X[c n2 + r]
        for (u32 r = threadIdx.x; r < n2; r += blockDim.x) {
            u64 si = 0, so = 0;
            for (u32 cc = 0; cc < n1; cc++) { si = addm(si, a[cc * n2 + r], q); so = addm(so, b[cc * n2 + r], q); }
            s_in[r] = si; s_out[r] = so; wv[r] = w_b1[r] % q; wh[r] = wv[r];
        }
        __syncthreads();
        sub_ntts(wh, c, 1, lg2, pw, n1, -1, nullptr, 0, m);
        u64 l = 0, r = 0;
        for (u32 i = threadIdx.x; i < n2; i += blockDim.x) { l = addm(l, mm(wh[i], s_in[i], m), q); r = addm(r, mm(wv[i], s_out[i], m), q); }
        l = block_sum(l, q, red); r = block_sum(r, q, red);
        if (l != r) flag |= 1;
    }
    // ---- stage 2: twiddle T[c n2 + r] = wN^(r c), intra fold check, inter check ----
    {
        const u64 fmask = (1ull << fold_bits) - 1;
        u32 intra = 0;
        u64 l = 0, r = 0;
        for (u32 i = threadIdx.x; i < N; i += blockDim.x) {
            const u32 cc = i >> lg2, rr = i & (n2 - 1);
            const u64 T = pw[((size_t)rr * cc) & (N - 1)], pre = b[i], ph = phi[rr] % q;
            const u64 vt = mm(pre, T, m);
            u64 v = vt;
            if (nf) v = corrupt(v, s1 + i, f, nf, m);
            if (v != vt && fold_sum(vt, fold_bits, fmask) != fold_sum(v, fold_bits, fmask)) intra = 1;
            b[i] = v;
            l = addm(l, mm(ph, v, m), q);
            r = addm(r, mm(mm(ph, T, m), pre, m), q);
        }
        if (__syncthreads_or(intra)) flag |= 2;
        l = block_sum(l, q, red); r = block_sum(r, q, red);
        if (l != r) flag |= 4;
    }
    // ---- stage 3: transpose, n2 blocks of n1 (root w_n1 = wN^n2) ----
    for (u32 i = threadIdx.x; i < N; i += blockDim.x) {
        const u32 cc = i >> lg2, rr = i & (n2 - 1);
        const u64 v = b[i];
        a[rr * n1 + cc] = v;                  // C_in
    }
    __syncthreads();
    for (u32 i = threadIdx.x; i < N; i += blockDim.x) b[i] = a[i];
    __syncthreads();
    sub_ntts(b, c, n2, lg1, pw, n2, s3, f, nf, m);
    {   // batch check on rows: s[c] = sum_r X[r n1 + c]
        for (u32 cc = threadIdx.x; cc < n1; cc += blockDim.x) {
            u64 si = 0, so = 0;
            for (u32 rr = 0; rr < n2; rr++) { si = addm(si, a[rr * n1 + cc], q); so = addm(so, b[rr * n1 + cc], q); }
            s_in[cc] = si; s_out[cc] = so; wv[cc] = w_b2[cc] % q; wh[cc] = wv[cc];
        }
        __syncthreads();
        sub_ntts(wh, c, 1, lg1, pw, n2, -1, nullptr, 0, m);
        u64 l = 0, r = 0;
        for (u32 i = threadIdx.x; i < n1; i += blockDim.x) { l = addm(l, mm(wh[i], s_in[i], m), q); r = addm(r, mm(wv[i], s_out[i], m), q); }
        l = block_sum(l, q, red); r = block_sum(r, q, red);
        if (l != r) flag |= 8;
    }
    if (out)
        for (u32 i = threadIdx.x; i < N; i += blockDim.x) out[(size_t)tr * N + i] = b[i];
    if (threadIdx.x == 0) flags[tr] = flag;
}

// ---- the baseline the reference compared against: standard iterative NTT + lazy bucket check + global check
// (rfhe_framewk/src/ntt_test/non_four_step_ntt_sim.cpp) ---------------------------------------------------------------
// ntt_standard_lazy (:164-218): one length-N radix-2 DIT transform whose every product v = A[k+half] * w is handed to the
// LazyChecker (:135-158) twice -- the value the butterfly consumes (after injection) into sum_out[op % buckets], the
// UNREDUCED 128-bit product into sum_ref[op % buckets] -- and the check is sum_out == sum_ref (mod Q) per bucket: a
// reduction-unit fault shows up in its bucket.  Device form: the sums are exact integers kept as sums of 32-bit pieces (two of
// v, four of the product; each < 2^32 * N/2 log N / buckets), reduced mod Q once at the end.  Operation numbering as in
// sub_ntts (stage-major, then group, then k: the simulator's loop order).
// lz: [buckets][6] accumulators, pieces 0-1 of v, 2-5 of the product.
__device__ void std_ntt_lazy(u64* x, u64* tmp, int lg, const u64* pw, const ReliaFault* f, u32 nf, const B200Mod& m, u64* lz, u32 buckets) {
    const u32 len = 1u << lg;
    for (u32 i = threadIdx.x; i < len; i += blockDim.x) tmp[brev(i, lg)] = x[i];
    __syncthreads();
    for (u32 i = threadIdx.x; i < len; i += blockDim.x) x[i] = tmp[i];
    __syncthreads();
    const u32 hb = len >> 1;
    for (int s = 0; s < lg; s++) {
        const u32 half = 1u << s;
        for (u32 b = threadIdx.x; b < hb; b += blockDim.x) {
            const u32 grp = b >> s, k = b & (half - 1);
            u64* p = x + (grp << (s + 1)) + k;
            const u64 w = pw[(size_t)k << (lg - 1 - s)];
            const u64 u = p[0], t = p[half];
            u64 v = mm(t, w, m);
            const long long op = (long long)s * hb + b;
            if (nf) v = corrupt(v, op, f, nf, m);
            if (lz) {
                const u64 lo = t * w, hi = __umul64hi(t, w);
                u64* a = lz + (size_t)(op % buckets) * 6;
                atomicAdd((unsigned long long*)a + 0, (unsigned long long)(v & 0xffffffffull));
                atomicAdd((unsigned long long*)a + 1, (unsigned long long)(v >> 32));
                atomicAdd((unsigned long long*)a + 2, (unsigned long long)(lo & 0xffffffffull));
                atomicAdd((unsigned long long*)a + 3, (unsigned long long)(lo >> 32));
                atomicAdd((unsigned long long*)a + 4, (unsigned long long)(hi & 0xffffffffull));
                atomicAdd((unsigned long long*)a + 5, (unsigned long long)(hi >> 32));
            }
            p[0] = addm(u, v, m.q);
            p[half] = addm(u, m.q - v, m.q);
        }
        __syncthreads();
    }
}

// sm: pw[N] | a[N] input | b[N] transform | c[N] scratch | wv[N] | wh[N] | lz[buckets][6] | red[32]
__global__ void __launch_bounds__(256) k_relia_standard(const u64* __restrict__ in, u64* __restrict__ out, const u64* __restrict__ wts, u32 N, int lg,
                                                        B200Mod m, const u64* __restrict__ pw_g, u32 buckets, const ReliaFault* faults,
                                                        const u32* fault_begin, u32* flags) {
    extern __shared__ u64 sm[];
    u64 *pw = sm, *a = sm + N, *b = a + N, *c = b + N, *wv = c + N, *wh = wv + N, *lz = wh + N, *red = lz + (size_t)buckets * 6;
    const u32 tr = blockIdx.x;
    const u64 q = m.q;
    const ReliaFault* f = faults + fault_begin[tr];
    const u32 nf = fault_begin[tr + 1] - fault_begin[tr];
    for (u32 i = threadIdx.x; i < N; i += blockDim.x) {
        pw[i] = pw_g[i];
        const u64 v = in[(size_t)tr * N + i] % q;
        a[i] = v; b[i] = v;
        const u64 w = wts[(size_t)tr * N + i] % q;
        wv[i] = w; wh[i] = w;
    }
    for (u32 i = threadIdx.x; i < buckets * 6; i += blockDim.x) lz[i] = 0;
    __syncthreads();
    std_ntt_lazy(b, c, lg, pw, f, nf, m, lz, buckets);
    u32 flag = 0;
    {   // LazyChecker::verify (:151-157)
        u32 bad = 0;
        const u64 r32 = ((u64)1 << 32) % q, r64 = mm(r32, r32, m), r96 = mm(r64, r32, m);
        for (u32 i = threadIdx.x; i < buckets; i += blockDim.x) {
            const u64* e = lz + (size_t)i * 6;
            const u64 so = addm(e[0] % q, mm(e[1] % q, r32, m), q);
            u64 sr = addm(e[2] % q, mm(e[3] % q, r32, m), q);
            sr = addm(sr, mm(e[4] % q, r64, m), q);
            sr = addm(sr, mm(e[5] % q, r96, m), q);
            if (so != sr) bad = 1;
        }
        if (__syncthreads_or(bad)) flag |= 1;
    }
    // check_abft_global (:220-235): NTT(w) . input == w . output
    std_ntt_lazy(wh, c, lg, pw, nullptr, 0, m, nullptr, 1);
    u64 l = 0, r = 0;
    for (u32 i = threadIdx.x; i < N; i += blockDim.x) { l = addm(l, mm(wh[i], a[i], m), q); r = addm(r, mm(wv[i], b[i], m), q); }
    l = block_sum(l, q, red); r = block_sum(r, q, red);
    if (l != r) flag |= 2;
    if (out)
        for (u32 i = threadIdx.x; i < N; i += blockDim.x) out[(size_t)tr * N + i] = b[i];
    if (threadIdx.x == 0) flags[tr] = flag;
}

}  // namespace b200

using namespace b200;

extern "C" int b200fhe_relia_op_ranges(size_t N, uint64_t* s1_end, uint64_t* s2_end, uint64_t* s3_end) {
    int lg = 0;
    while (((size_t)1 << lg) < N) lg++;
    if (((size_t)1 << lg) != N || N < 4 || (lg & 1)) return b200_set_err(B200FHE_EINVAL, "relia: N must be an even power of two >= 4");
    const int lg1 = lg / 2, lg2 = lg - lg1;           // n1 = round(sqrt N), n2 = N / n1 (:296-297): a power of two only for even log N
    const uint64_t n1 = (uint64_t)1 << lg1, n2 = (uint64_t)1 << lg2;
    const uint64_t a = n1 * (n2 / 2) * lg2, b = a + N, c = b + n2 * (n1 / 2) * lg1;
    if (s1_end) *s1_end = a;
    if (s2_end) *s2_end = b;
    if (s3_end) *s3_end = c;
    return 0;
}

extern "C" int b200fhe_relia_four_step_host(uint64_t q, uint64_t g, size_t N, int fold_bits, const uint64_t* h_in, uint64_t* h_out,
                                            const uint64_t* h_weights, size_t trials, const b200fhe_relia_fault* faults, size_t n_faults,
                                            uint32_t* h_flags, int device) {
    if (!h_in || !h_weights || !h_flags || (n_faults && !faults)) return b200_set_err(B200FHE_EINVAL, "relia: null argument");
    int lg = 0;
    while (((size_t)1 << lg) < N) lg++;
    if (((size_t)1 << lg) != N || N < 4 || N > 4096 || (lg & 1))
        return b200_set_err(B200FHE_EUNSUPPORTED, "relia: N=%zu must be an even power of two in [4, 4096] (the simulator's n1 = round(sqrt N))", N);
    if (q < 3 || (q >> 60) || !nt::is_prime(q) || (q - 1) % N) return b200_set_err(B200FHE_EINVAL, "relia: modulus must be a prime < 2^60 with N | q-1");
    if (fold_bits < 1 || fold_bits > 32) return b200_set_err(B200FHE_EINVAL, "relia: fold_bits out of range");
    if (trials == 0) return 0;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return b200_set_err(B200FHE_ECUDA, "no CUDA device: this library has no CPU fallback");
    if (device < 0 || device >= ndev) return b200_set_err(B200FHE_EINVAL, "device %d of %d", device, ndev);
    cudaSetDevice(device);
    const int lg1 = lg / 2, lg2 = lg - lg1;
    const size_t n1 = (size_t)1 << lg1, n2 = (size_t)1 << lg2, nm = n1 > n2 ? n1 : n2;
    const uint64_t wN = nt::powmod(g, (q - 1) / N, q);           // root_of_unity(N) (:95-97)
    if (nt::powmod(wN, N / 2, q) != q - 1) return b200_set_err(B200FHE_EINVAL, "relia: g=%llu does not give a primitive N-th root", (unsigned long long)g);
    std::vector<uint64_t> pw(N);
    pw[0] = 1;
    for (size_t i = 1; i < N; i++) pw[i] = nt::mulmod(pw[i - 1], wN, q);
    uint64_t total = 0;
    b200fhe_relia_op_ranges(N, nullptr, nullptr, &total);
    // faults grouped by trial (stable: entries of one trial keep their order, the simulator applies a sorted plan in order)
    std::vector<uint32_t> begin(trials + 1, 0);
    for (size_t i = 0; i < n_faults; i++) {
        if (faults[i].trial >= trials || faults[i].op >= total || faults[i].kind < 1 || faults[i].kind > 2)
            return b200_set_err(B200FHE_EINVAL, "relia: fault %zu out of range", i);
        begin[faults[i].trial + 1]++;
    }
    for (size_t t = 0; t < trials; t++) begin[t + 1] += begin[t];
    std::vector<ReliaFault> fl(n_faults ? n_faults : 1);
    {
        std::vector<uint32_t> pos(begin.begin(), begin.end() - 1);
        for (size_t i = 0; i < n_faults; i++) fl[pos[faults[i].trial]++] = ReliaFault{faults[i].op, faults[i].payload, faults[i].kind, 0};
    }
    const B200Mod m = b200_make_mod(q, 0, 1);
    const size_t wper = 2 * n2 + n1;
    u64 *d_in = nullptr, *d_out = nullptr, *d_w = nullptr, *d_pw = nullptr;
    u32 *d_begin = nullptr, *d_flags = nullptr;
    ReliaFault* d_f = nullptr;
    cudaError_t e = cudaMalloc(&d_in, trials * N * 8);
    if (e == cudaSuccess && h_out) e = cudaMalloc(&d_out, trials * N * 8);
    if (e == cudaSuccess) e = cudaMalloc(&d_w, trials * wper * 8);
    if (e == cudaSuccess) e = cudaMalloc(&d_pw, N * 8);
    if (e == cudaSuccess) e = cudaMalloc(&d_begin, (trials + 1) * 4);
    if (e == cudaSuccess) e = cudaMalloc(&d_flags, trials * 4);
    if (e == cudaSuccess) e = cudaMalloc(&d_f, sizeof(ReliaFault) * fl.size());
    if (e == cudaSuccess) e = cudaMemcpy(d_in, h_in, trials * N * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_w, h_weights, trials * wper * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_pw, pw.data(), N * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_begin, begin.data(), (trials + 1) * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_f, fl.data(), sizeof(ReliaFault) * fl.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        const size_t smem = (4 * N + 4 * nm + 32) * 8;
        e = set_smem_once<k_relia_four_step>((4 * 4096 + 4 * 64 + 32) * 8);
        if (e == cudaSuccess) {
            k_relia_four_step<<<(unsigned)trials, 256, smem>>>(d_in, d_out, d_w, (u32)N, lg1, lg2, m, d_pw, fold_bits, d_f, d_begin, d_flags);
            b200_count_launches(1);
            e = cudaGetLastError();
        }
    }
    if (e == cudaSuccess && h_out) e = cudaMemcpy(h_out, d_out, trials * N * 8, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(h_flags, d_flags, trials * 4, cudaMemcpyDeviceToHost);
    cudaFree(d_in); cudaFree(d_out); cudaFree(d_w); cudaFree(d_pw); cudaFree(d_begin); cudaFree(d_flags); cudaFree(d_f);
    if (e != cudaSuccess) return b200_set_err(B200FHE_ECUDA, "relia: %s", cudaGetErrorString(e));
    return 0;
}

extern "C" int b200fhe_relia_standard_host(uint64_t q, uint64_t g, size_t N, const uint64_t* h_in, uint64_t* h_out, const uint64_t* h_weights,
                                           size_t trials, const b200fhe_relia_fault* faults, size_t n_faults, uint32_t* h_flags, int device) {
    if (!h_in || !h_weights || !h_flags || (n_faults && !faults)) return b200_set_err(B200FHE_EINVAL, "relia_standard: null argument");
    int lg = 0;
    while (((size_t)1 << lg) < N) lg++;
    if (((size_t)1 << lg) != N || N < 2 || N > 4096) return b200_set_err(B200FHE_EUNSUPPORTED, "relia_standard: N=%zu must be a power of two in [2, 4096]", N);
    if (q < 3 || (q >> 60) || !nt::is_prime(q) || (q - 1) % N)
        return b200_set_err(B200FHE_EINVAL, "relia_standard: modulus must be a prime < 2^60 with N | q-1");
    if (trials == 0) return 0;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return b200_set_err(B200FHE_ECUDA, "no CUDA device: this library has no CPU fallback");
    if (device < 0 || device >= ndev) return b200_set_err(B200FHE_EINVAL, "device %d of %d", device, ndev);
    cudaSetDevice(device);
    const uint64_t wN = nt::powmod(g, (q - 1) / N, q);
    if (nt::powmod(wN, N / 2, q) != q - 1) return b200_set_err(B200FHE_EINVAL, "relia_standard: g=%llu does not give a primitive N-th root", (unsigned long long)g);
    std::vector<uint64_t> pw(N);
    pw[0] = 1;
    for (size_t i = 1; i < N; i++) pw[i] = nt::mulmod(pw[i - 1], wN, q);
    const uint64_t total = (uint64_t)(N / 2) * lg;                 // the simulator's dry-run op count (:262-265)
    uint32_t buckets = 1;
    while ((uint64_t)(buckets + 1) * (buckets + 1) <= N) buckets++;   // (int)sqrt(n) (:170)
    std::vector<uint32_t> begin(trials + 1, 0);
    for (size_t i = 0; i < n_faults; i++) {
        if (faults[i].trial >= trials || faults[i].op >= total || faults[i].kind < 1 || faults[i].kind > 2)
            return b200_set_err(B200FHE_EINVAL, "relia_standard: fault %zu out of range", i);
        begin[faults[i].trial + 1]++;
    }
    for (size_t t = 0; t < trials; t++) begin[t + 1] += begin[t];
    std::vector<ReliaFault> fl(n_faults ? n_faults : 1);
    {
        std::vector<uint32_t> pos(begin.begin(), begin.end() - 1);
        for (size_t i = 0; i < n_faults; i++) fl[pos[faults[i].trial]++] = ReliaFault{faults[i].op, faults[i].payload, faults[i].kind, 0};
    }
    const B200Mod m = b200_make_mod(q, 0, 1);
    u64 *d_in = nullptr, *d_out = nullptr, *d_w = nullptr, *d_pw = nullptr;
    u32 *d_begin = nullptr, *d_flags = nullptr;
    ReliaFault* d_f = nullptr;
    cudaError_t e = cudaMalloc(&d_in, trials * N * 8);
    if (e == cudaSuccess && h_out) e = cudaMalloc(&d_out, trials * N * 8);
    if (e == cudaSuccess) e = cudaMalloc(&d_w, trials * N * 8);
    if (e == cudaSuccess) e = cudaMalloc(&d_pw, N * 8);
    if (e == cudaSuccess) e = cudaMalloc(&d_begin, (trials + 1) * 4);
    if (e == cudaSuccess) e = cudaMalloc(&d_flags, trials * 4);
    if (e == cudaSuccess) e = cudaMalloc(&d_f, sizeof(ReliaFault) * fl.size());
    if (e == cudaSuccess) e = cudaMemcpy(d_in, h_in, trials * N * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_w, h_weights, trials * N * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_pw, pw.data(), N * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_begin, begin.data(), (trials + 1) * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_f, fl.data(), sizeof(ReliaFault) * fl.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        const size_t smem = (6 * N + (size_t)buckets * 6 + 32) * 8;
        e = set_smem_once<k_relia_standard>((6 * 4096 + 64 * 6 + 32) * 8);
        if (e == cudaSuccess) {
            k_relia_standard<<<(unsigned)trials, 256, smem>>>(d_in, d_out, d_w, (u32)N, lg, m, d_pw, buckets, d_f, d_begin, d_flags);
            b200_count_launches(1);
            e = cudaGetLastError();
        }
    }
    if (e == cudaSuccess && h_out) e = cudaMemcpy(h_out, d_out, trials * N * 8, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(h_flags, d_flags, trials * 4, cudaMemcpyDeviceToHost);
    cudaFree(d_in); cudaFree(d_out); cudaFree(d_w); cudaFree(d_pw); cudaFree(d_begin); cudaFree(d_flags); cudaFree(d_f);
    if (e != cudaSuccess) return b200_set_err(B200FHE_ECUDA, "relia_standard: %s", cudaGetErrorString(e));
    return 0;
}
