// The protected four-step NTT exactly as prototyped in reliability_test/four_step_ntt_prot.py:
// cyclic transform of N = n*n points modulo any NTT prime (the prototype uses 998244353, g = 3,
// :16-17), computed as   Pi_k (I (x) W_n) S D (I (x) W_n) Pi_t   (:133-181)  with the prototype's
// scalar protection equality  sum(C) == colsum(A) . rowsum(B)  (:185-194) evaluated on BOTH
// matrix stages inside the kernel that computes them (:196-252 four_step_with_protection_vector):
//   stage 1: lhs1 = sum of all step-1 outputs,  rhs1 = sum_s colsum(I (x) W_n)[s] * x1[s]
//   stage 2: lhs2 = sum of all step-3 outputs,  rhs2 = sum_s colsum(I (x) W_n)[s] * x4[s]
// One CTA per transform; the vector, its intermediate and the power table w^e live in shared
// memory (N <= 4096, the sizes the O(N^2)-memory prototype can run).  Faults are injected where
// demo() injects them (:271-274: `C1[i] += delta` after the stage-1 multiply), or after stage 2.
// The production-size protected transform (N = 2^12..2^17, negacyclic, SEAL order) is the fused
// weighted-checksum path of ntt_kernels.cu; this kernel is the literal prototype on the GPU.
#include "../../include/b200fhe.h"
#include "ctx.hpp"
#include "numth.hpp"
#include <vector>

int b200_set_err(int code, const char* fmt, ...);
B200Mod b200_make_mod(uint64_t q, uint64_t itw1, uint64_t n_inv);

namespace b200 {

struct FsInject {
    u32 transform, stage;   // stage 1: after the first matrix stage, 2: after the second
    u32 row, pad;           // index s into the stage's output vector (prototype layout)
    u64 delta;
};

// 128-bit lazy accumulation of x*y (both < q < 2^60, at most 64 terms: < 2^126)
struct Acc128 {
    u64 lo = 0, hi = 0;
    __device__ __forceinline__ void mad(u64 x, u64 y) {
        const u64 l = x * y, h = __umul64hi(x, y);
        lo += l;
        hi += h + (lo < l);
    }
};

__device__ __forceinline__ u64 block_sum_mod(u64 v, u64 q, u64* red) {
    // v < q per thread; tree over the CTA with conditional subtraction
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int o = 16; o; o >>= 1) {
        u64 t = __shfl_down_sync(0xffffffffu, v, o);
        v += t;
        if (v >= q) v -= q;
    }
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    u64 r = 0;
    for (int i = 0; i < (int)(blockDim.x >> 5); i++) {
        r += red[i];
        if (r >= q) r -= q;
    }
    return r;
}

// sm layout: pw[N] (w^e) | x[N] | y[N] | cs[n] | red[32]
__global__ void __launch_bounds__(256) k_four_step_prot(const u64* __restrict__ in, u64* __restrict__ out, u32 N, u32 n, B200Mod m,
                                                        const u64* __restrict__ pw_g, const u64* __restrict__ cs_g, u64* checks,
                                                        u32* verdict, const FsInject* inj, u32 n_inj) {
    extern __shared__ u64 sm[];
    u64 *pw = sm, *x = sm + N, *y = x + N, *cs = y + N, *red = cs + n;
    const u32 tr = blockIdx.x;
    const u64 q = m.q;
    const u64* a = in + (size_t)tr * N;
    for (u32 i = threadIdx.x; i < N; i += blockDim.x) pw[i] = pw_g[i];
    for (u32 i = threadIdx.x; i < n; i += blockDim.x) cs[i] = cs_g[i];
    // Pi_t: x1[t1*n + t2] = a[t1 + n*t2]  (four_step_ntt_prot.py:147-152), inputs taken mod q
    for (u32 s = threadIdx.x; s < N; s += blockDim.x) {
        const u32 t1 = s / n, t2 = s % n;
        x[s] = a[t1 + n * t2] % q;
    }
    __syncthreads();
    u64 sums[4];
    for (int stage = 1; stage <= 2; stage++) {
        // rhs = colsum(I (x) W_n) . input ;  colsum[s] = cs[s mod n]
        u64 part = 0;
        for (u32 s = threadIdx.x; s < N; s += blockDim.x) {
            part += mulmod_barrett(cs[s % n], x[s], m);
            if (part >= q) part -= q;
        }
        const u64 rhs = block_sum_mod(part, q, red);
        // (I (x) W_n): y[b*n + k] = sum_t W_n[k][t] x[b*n + t],  W_n[k][t] = (w^n)^(k t) = pw[n * (k t mod n)]
        for (u32 s = threadIdx.x; s < N; s += blockDim.x) {
            const u32 b = s / n, k = s % n;
            Acc128 acc;
            u32 e = 0;                                   // k*t mod n
            for (u32 t = 0; t < n; t++) {
                acc.mad(x[b * n + t], pw[n * e]);
                e += k;
                if (e >= n) e -= n;
            }
            y[s] = barrett128(acc.lo, acc.hi, m);
        }
        __syncthreads();
        for (u32 i = threadIdx.x; i < n_inj; i += blockDim.x)       // demo(): C1_fault[i] = C1[i] + delta
            if (inj[i].transform == tr && inj[i].stage == (u32)stage && inj[i].row < N) {
                u64 v = y[inj[i].row] + inj[i].delta % q;
                y[inj[i].row] = v >= q ? v - q : v;
            }
        __syncthreads();
        part = 0;
        for (u32 s = threadIdx.x; s < N; s += blockDim.x) {
            part += y[s];
            if (part >= q) part -= q;
        }
        const u64 lhs = block_sum_mod(part, q, red);
        sums[2 * (stage - 1)] = lhs;
        sums[2 * (stage - 1) + 1] = rhs;
        if (stage == 1) {
            // D then S (:154-170): x4[k2*n + t1] = y[t1*n + k2] * w^(t1 k2)
            for (u32 s = threadIdx.x; s < N; s += blockDim.x) {
                const u32 k2 = s / n, t1 = s % n;
                x[s] = mulmod_barrett(y[t1 * n + k2], pw[t1 * k2], m);
            }
            __syncthreads();
        }
    }
    // Pi_k (:175-180): out[k1*n + k2] = y[k2*n + k1]
    for (u32 k = threadIdx.x; k < N; k += blockDim.x) out[(size_t)tr * N + k] = y[(k % n) * n + k / n];
    if (threadIdx.x == 0) {
        for (int i = 0; i < 4; i++) checks[(size_t)tr * 4 + i] = sums[i];
        verdict[tr] = (sums[0] != sums[1] ? 1u : 0u) | (sums[2] != sums[3] ? 2u : 0u);
    }
}

}  // namespace b200

using namespace b200;

extern "C" int b200fhe_four_step_protected_host(uint64_t mod, uint64_t g, size_t N, const uint64_t* h_in, uint64_t* h_out,
                                                size_t batch, uint64_t* h_checks, uint32_t* h_verdict,
                                                const b200fhe_fs_inject* plan, size_t n_plan, int device) {
    if (!h_in || !h_out || !h_checks || !h_verdict || (n_plan && !plan)) return b200_set_err(B200FHE_EINVAL, "four_step: null argument");
    size_t n = 1;
    while (n * n < N) n++;
    if (n * n != N || N < 4 || N > 4096) return b200_set_err(B200FHE_EUNSUPPORTED, "four_step: N=%zu must be a perfect square in [4, 4096]", N);
    if (mod < 3 || (mod >> 60) || !nt::is_prime(mod) || (mod - 1) % N) return b200_set_err(B200FHE_EINVAL, "four_step: modulus must be a prime < 2^60 with N | mod-1");
    if (n_plan > (size_t)B200FHE_MAX_INJECT * (batch ? batch : 1)) return b200_set_err(B200FHE_EINVAL, "four_step: too many injections (at most %d per transform of the batch)", B200FHE_MAX_INJECT);
    if (batch == 0) return 0;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return b200_set_err(B200FHE_ECUDA, "no CUDA device: this library has no CPU fallback");
    if (device < 0 || device >= ndev) return b200_set_err(B200FHE_EINVAL, "device %d of %d", device, ndev);
    cudaSetDevice(device);
    const uint64_t w = nt::powmod(g, (mod - 1) / N, mod);        // root_of_unity (:24-25)
    if (nt::powmod(w, N / 2, mod) != mod - 1) return b200_set_err(B200FHE_EINVAL, "four_step: g=%llu does not give a primitive N-th root", (unsigned long long)g);
    std::vector<uint64_t> pw(N), cs(n, 0);
    pw[0] = 1;
    for (size_t i = 1; i < N; i++) pw[i] = nt::mulmod(pw[i - 1], w, mod);
    for (size_t t = 0; t < n; t++)                               // col_sums of W_n (:30-32): sum_k (w^n)^(k t)
        for (size_t k = 0; k < n; k++) cs[t] = (cs[t] + pw[n * ((k * t) % n)]) % mod;
    std::vector<FsInject> inj(n_plan ? n_plan : 1);
    for (size_t i = 0; i < n_plan; i++) {
        if (plan[i].transform >= batch || plan[i].stage < 1 || plan[i].stage > 2 || plan[i].row >= N)
            return b200_set_err(B200FHE_EINVAL, "four_step: injection %zu out of range", i);
        inj[i] = FsInject{plan[i].transform, plan[i].stage, (u32)plan[i].row, 0, plan[i].delta};
    }
    const B200Mod m = b200_make_mod(mod, 0, 1);
    u64 *d_in = nullptr, *d_out = nullptr, *d_pw = nullptr, *d_cs = nullptr, *d_chk = nullptr;
    u32* d_ver = nullptr;
    FsInject* d_inj = nullptr;
    cudaError_t e = cudaMalloc(&d_in, batch * N * 8);
    if (e == cudaSuccess) e = cudaMalloc(&d_out, batch * N * 8);
    if (e == cudaSuccess) e = cudaMalloc(&d_pw, N * 8);
    if (e == cudaSuccess) e = cudaMalloc(&d_cs, n * 8);
    if (e == cudaSuccess) e = cudaMalloc(&d_chk, batch * 32);
    if (e == cudaSuccess) e = cudaMalloc(&d_ver, batch * 4);
    if (e == cudaSuccess) e = cudaMalloc(&d_inj, sizeof(FsInject) * inj.size());
    if (e == cudaSuccess) e = cudaMemcpy(d_in, h_in, batch * N * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_pw, pw.data(), N * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_cs, cs.data(), n * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_inj, inj.data(), sizeof(FsInject) * inj.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        const size_t smem = (3 * N + n + 32) * 8;
        cudaFuncSetAttribute(k_four_step_prot, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_four_step_prot<<<(unsigned)batch, 256, smem>>>(d_in, d_out, (u32)N, (u32)n, m, d_pw, d_cs, d_chk, d_ver, d_inj, (u32)n_plan);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(h_out, d_out, batch * N * 8, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(h_checks, d_chk, batch * 32, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(h_verdict, d_ver, batch * 4, cudaMemcpyDeviceToHost);
    cudaFree(d_in); cudaFree(d_out); cudaFree(d_pw); cudaFree(d_cs); cudaFree(d_chk); cudaFree(d_ver); cudaFree(d_inj);
    if (e != cudaSuccess) return b200_set_err(B200FHE_ECUDA, "four_step: %s", cudaGetErrorString(e));
    return 0;
}
