// Evaluator steps either side of the key-switch (device side): the element-wise and permutation kernels
// that phantom::multiply / add_inplace / rotate_inplace run around it in reliability_test/dotprod_test.cu:113-152.
// All operate on NTT-form ciphertext polynomials laid out [n_poly][L][N] (dotprod_test.cu:42-46) and are HBM-bound
// streaming kernels: one thread = two consecutive coefficients, every word read and written once.
#include "../../include/b200fhe.h"
#include "ctx.hpp"

using namespace b200;

int b200_set_err(int code, const char* fmt, ...);
void b200_count_launches(unsigned n);

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return b200_set_err(B200FHE_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                                __FILE__, __LINE__);                                               \
    } while (0)

namespace b200 {

// (d0, d1, d2) = (a0 b0, a0 b1 + a1 b0, a1 b1) mod q_l; inputs any 64-bit words (taken mod q), outputs in [0,q)
__global__ void __launch_bounds__(256) k_ct_tensor(const u64* __restrict__ a, const u64* __restrict__ b, u64* __restrict__ out,
                                                   const B200Mod* mods, u32 N, u32 L, u32 start_mod) {
    const u32 limb = blockIdx.y;
    const B200Mod m = mods[start_mod + limb];
    const size_t i2 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i2 * 2 >= N) return;
    const size_t off = ((size_t)limb * N) / 2 + i2, poly2 = ((size_t)L * N) / 2;
    const ulonglong2 a0 = __ldcs(reinterpret_cast<const ulonglong2*>(a) + off), a1 = __ldcs(reinterpret_cast<const ulonglong2*>(a) + poly2 + off);
    const ulonglong2 b0 = __ldcs(reinterpret_cast<const ulonglong2*>(b) + off), b1 = __ldcs(reinterpret_cast<const ulonglong2*>(b) + poly2 + off);
    // reduce128: three folds of 2^bits = delta for the pseudo-Mersenne primes and in-range operands (the usual case: 5 IMAD.WIDE
    // instead of a 128-bit Barrett's ~17), Barrett otherwise.  Measured: 40 -> 37 us at C2's shape, i.e. the multiplier was not the
    // bound; neither is memory-level parallelism (2 / 4 pairs per thread with all loads up front: 39 / 46 us,
    // profiles/r3e_ct_tensor_sweep.txt) -- a 37 us kernel of 3.9 waves pays its ramp and tail
    auto mul = [&](u64 x, u64 y) { return reduce128(x * y, __umul64hi(x, y), m); };
    auto mul2 = [&](u64 x0, u64 y0, u64 x1, u64 y1) {          // x0 y0 + x1 y1 mod q with one reduction
        const u64 l0 = x0 * y0, h0 = __umul64hi(x0, y0), l1 = x1 * y1, h1 = __umul64hi(x1, y1);
        const u64 lo = l0 + l1, c = lo < l0;
        const u64 t = h0 + h1, hi = t + c;
        const u64 top = (u64)(t < h0) + (u64)(hi < t);         // carry out of the high word (0 or 1)
        if (top == 0) return reduce128(lo, hi, m);             // operands below q: the sum is below 2 q^2
        const u64 r1 = barrett128(hi, top, m);                 // (top:hi) mod q
        return barrett128(lo, r1, m);
    };
    ulonglong2 d0, d1, d2;
    d0.x = mul(a0.x, b0.x); d0.y = mul(a0.y, b0.y);
    d2.x = mul(a1.x, b1.x); d2.y = mul(a1.y, b1.y);
    d1.x = mul2(a0.x, b1.x, a1.x, b0.x); d1.y = mul2(a0.y, b1.y, a1.y, b0.y);
    reinterpret_cast<ulonglong2*>(out)[off] = d0;
    reinterpret_cast<ulonglong2*>(out)[poly2 + off] = d1;
    reinterpret_cast<ulonglong2*>(out)[2 * poly2 + off] = d2;
}

// a = (a + b) mod q (SUB = false) or (a - b) mod q; inputs in [0,q)
template <bool SUB>
__global__ void __launch_bounds__(256) k_poly_addsub(u64* __restrict__ a, const u64* __restrict__ b, const B200Mod* mods, u32 N, u32 L,
                                                     u32 start_mod) {
    const u32 limb = blockIdx.y;
    const u64 q = mods[start_mod + limb].q;
    const size_t i2 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i2 * 2 >= N) return;
    const size_t off = (((size_t)blockIdx.z * L + limb) * N) / 2 + i2;
    ulonglong2 x = reinterpret_cast<const ulonglong2*>(a)[off];
    const ulonglong2 y = __ldcs(reinterpret_cast<const ulonglong2*>(b) + off);
    if (SUB) {
        x.x = x.x - y.x + (x.x < y.x ? q : 0);
        x.y = x.y - y.y + (x.y < y.y ? q : 0);
    } else {
        x.x += y.x; x.x -= (x.x >= q ? q : 0);
        x.y += y.y; x.y -= (x.y >= q ? q : 0);
    }
    reinterpret_cast<ulonglong2*>(a)[off] = x;
}

// Galois automorphism x -> x^g on an NTT-form polynomial (bit-reversed order): the value at psi^(2k+1) of a(x^g) is the
// value of a at psi^(g (2k+1)), so the automorphism is the index permutation k -> ((g (2k+1) mod 2N) - 1) / 2 and needs
// no sign changes (SEAL GaloisTool::apply_galois_ntt).  Writes are coalesced, reads are the gather.
__global__ void __launch_bounds__(256) k_galois_ntt(const u64* __restrict__ in, u64* __restrict__ out, u32 N, int logn, u32 galois) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const size_t base = (size_t)blockIdx.y * N;
    const u32 k = __brev(i) >> (32 - logn);
    const u32 e = (u32)(((u64)galois * (2 * k + 1)) & (2 * (u64)N - 1));
    const u32 src = __brev((e - 1) >> 1) >> (32 - logn);
    out[base + i] = __ldg(in + base + src);
}

}  // namespace b200

static int check_range(const b200fhe_ctx* c, size_t L, size_t start, const char* what) {
    if (!c) return b200_set_err(B200FHE_EINVAL, "%s: null context", what);
    if (start + L > (size_t)c->L) return b200_set_err(B200FHE_EINVAL, "%s: limbs [%zu, %zu) exceed the context's %d moduli", what, start, start + L, c->L);
    for (size_t i = start; i < start + L; i++)
        if (!c->moduli[i]) return b200_set_err(B200FHE_EINVAL, "%s: modulus %zu not set", what, i);
    if (c->N < 2) return b200_set_err(B200FHE_EINVAL, "%s: N too small", what);
    return 0;
}

extern "C" int b200fhe_ct_tensor(b200fhe_ctx* c, const uint64_t* d_a, const uint64_t* d_b, uint64_t* d_out, size_t L, size_t start_mod,
                                 void* stream) {
    int rc = check_range(c, L, start_mod, "ct_tensor");
    if (rc) return rc;
    if (L == 0) return 0;
    if (!d_a || !d_b || !d_out) return b200_set_err(B200FHE_EINVAL, "ct_tensor: null argument");
    CK(cudaSetDevice(c->device));
    dim3 grid((unsigned)((c->N / 2 + 255) / 256), (unsigned)L);
    k_ct_tensor<<<grid, 256, 0, (cudaStream_t)stream>>>((const u64*)d_a, (const u64*)d_b, (u64*)d_out, c->d_mods, (u32)c->N, (u32)L, (u32)start_mod);
    b200_count_launches(1);
    CK(cudaGetLastError());
    return 0;
}

static int addsub(b200fhe_ctx* c, uint64_t* d_a, const uint64_t* d_b, size_t n_poly, size_t L, size_t start_mod, bool sub, void* stream) {
    int rc = check_range(c, L, start_mod, sub ? "poly_sub" : "poly_add");
    if (rc) return rc;
    if (L == 0 || n_poly == 0) return 0;
    if (!d_a || !d_b || n_poly > 65535) return b200_set_err(B200FHE_EINVAL, "poly_add/sub: bad argument");
    CK(cudaSetDevice(c->device));
    dim3 grid((unsigned)((c->N / 2 + 255) / 256), (unsigned)L, (unsigned)n_poly);
    if (sub) k_poly_addsub<true><<<grid, 256, 0, (cudaStream_t)stream>>>((u64*)d_a, (const u64*)d_b, c->d_mods, (u32)c->N, (u32)L, (u32)start_mod);
    else k_poly_addsub<false><<<grid, 256, 0, (cudaStream_t)stream>>>((u64*)d_a, (const u64*)d_b, c->d_mods, (u32)c->N, (u32)L, (u32)start_mod);
    b200_count_launches(1);
    CK(cudaGetLastError());
    return 0;
}
extern "C" int b200fhe_poly_add(b200fhe_ctx* c, uint64_t* d_a, const uint64_t* d_b, size_t n_poly, size_t L, size_t start_mod, void* stream) {
    return addsub(c, d_a, d_b, n_poly, L, start_mod, false, stream);
}
extern "C" int b200fhe_poly_sub(b200fhe_ctx* c, uint64_t* d_a, const uint64_t* d_b, size_t n_poly, size_t L, size_t start_mod, void* stream) {
    return addsub(c, d_a, d_b, n_poly, L, start_mod, true, stream);
}

extern "C" int b200fhe_galois_ntt(b200fhe_ctx* c, const uint64_t* d_in, uint64_t* d_out, uint32_t galois_elt, size_t n_limbs, void* stream) {
    if (!c || !d_in || !d_out) return b200_set_err(B200FHE_EINVAL, "galois_ntt: null argument");
    if (d_in == d_out) return b200_set_err(B200FHE_EINVAL, "galois_ntt: in place is not supported");
    if (!(galois_elt & 1) || galois_elt >= 2 * c->N) return b200_set_err(B200FHE_EINVAL, "galois_ntt: element %u is not odd in [1, 2N)", galois_elt);
    if (n_limbs == 0) return 0;
    if (n_limbs > 65535) return b200_set_err(B200FHE_EINVAL, "galois_ntt: too many limbs in one call");
    CK(cudaSetDevice(c->device));
    dim3 grid((unsigned)((c->N + 255) / 256), (unsigned)n_limbs);
    k_galois_ntt<<<grid, 256, 0, (cudaStream_t)stream>>>((const u64*)d_in, (u64*)d_out, (u32)c->N, c->log_n, galois_elt);
    b200_count_launches(1);
    CK(cudaGetLastError());
    return 0;
}
