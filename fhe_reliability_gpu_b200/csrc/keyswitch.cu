// Hybrid key-switch: the step relinearize_inplace / rotate_inplace spend their time in
// (call sites reliability_test/dotprod_test.cu:114,146; SEAL switch_key_inplace, callgrind.out.ckks:11081;
// NTT 48 % + base conversion 42 % of it at dnum = 3, profile_framewk/draw_dnum.py:5-9):
//
//   c  [n_q][N]  (NTT domain, moduli q_0..q_{n_q-1})
//   1. y = INTT(c) (D_d/q_j)^-1 mod q_j                  one batched launch over all digits; the conversion's scale step is
//                                                        folded into the transform's final N^-1 constant (d_mods_scaled)
//   2. per digit d (limbs [d a, (d+1) a)):
//      up_d[targets] = NTT( sum_j y_dj (D_d/q_j) mod target )                          (k_bconv_contract_tc + NTT)
//      up_d[digit limbs] = c[digit limbs]                                              (pass through)
//   3. acc[p] = sum_d up_d (.) evk_d[p],  p = 0,1, over all n_q + n_p limbs             (k_pmult, T = dnum)
//   4. ModDown: t = INTT(acc[p] mod P);  conv = NTT( BaseConv_{P->Q}(t) );
//      out[p] = (acc[p] - conv) P^-1 mod q_l                                           (k_moddown_combine)
//
// The context must hold the moduli in the order q_0..q_{n_q-1}, p_0..p_{n_p-1}.  evk is stored split into its Q and P
// limb ranges ([dnum][2][n_q][N] and [dnum][2][n_p][N]) so each range is one pmult launch.  Both base conversions are the
// HPS fast conversion (b200fhe_bconv): ModDown's result differs from the exact floor((acc - [acc]_P)/P) by at most n_p,
// the usual approximate-ModDown noise.  All work is enqueued on the caller's stream; nothing synchronises or allocates
// after create, so the whole switch can be captured in a CUDA graph.
#include "../../include/b200fhe.h"
#include "ctx.hpp"
#include "numth.hpp"
#include <vector>

#include <cstdlib>
using namespace b200;

int b200_set_err(int code, const char* fmt, ...);
void b200_count_launches(unsigned n);
int b200_nwt_backward_mods(b200fhe_ctx* c, uint64_t* d, size_t n_poly, size_t L, size_t start, const B200Mod* d_mods_alt, void* stream);
int b200_nwt_forward_map(b200fhe_ctx* c, uint64_t* d, const uint2* d_map, const uint2* h_map, size_t n, void* stream);
int b200_nwt_backward_mods_oop(b200fhe_ctx* c, const uint64_t* d_src, uint64_t* d, size_t n_poly, size_t L, size_t start, const B200Mod* d_mods_alt,
                               void* stream);
int b200_bconv_contract_skip(b200fhe_bconv* p, const uint64_t* d_y, uint64_t* d_out, size_t skip_at, size_t skip_rows, void* stream);

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return b200_set_err(B200FHE_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                                __FILE__, __LINE__);                                               \
    } while (0)
#define RC(call)            \
    do {                    \
        int rc_ = (call);   \
        if (rc_) return rc_; \
    } while (0)

struct b200fhe_keyswitch {
    b200fhe_ctx* ctx = nullptr;
    int n_q = 0, n_p = 0, dnum = 0, alpha = 0;
    std::vector<b200fhe_bconv*> up;      // per digit: digit limbs -> every other limb of Q u P
    b200fhe_bconv* down = nullptr;       // P -> Q
    ulonglong2* d_pinv = nullptr;        // [n_q] (P^-1 mod q_l, Shoup quotient)
    B200Mod* d_mods_scaled = nullptr;    // [n_q + n_p] the context's constants with N^-1 replaced by N^-1 * (conversion scale of the limb)
    // ModUp's converted limbs of ALL digits as one scatter list for the forward NTT ((d total + g, g) for every g outside digit d):
    // one launch pair of dnum (total - alpha) limbs instead of up to two short launches per digit (wave quantisation: 16- and
    // 32-limb launches fill 0.9 / 1.7 waves of the column pass)
    std::vector<uint2> up_map;
    uint2* d_up_map = nullptr;
    uint64_t* d_ws = nullptr;                 // workspace, see offsets
    size_t o_x = 0, o_y = 0, o_up = 0, o_accq = 0, o_accp = 0, o_conv = 0, ws_words = 0;
    // fault injection armed for the next protected apply (b200fhe_keyswitch_set_inject)
    int inj_conv = -1;
    std::vector<b200fhe_bconv_inject> inj;
};

namespace b200 {

// out[p][l][i] = (acc[p][l][i] - conv[p][l][i]) * Pinv_l mod q_l; one thread = two coefficients
__global__ void __launch_bounds__(256) k_moddown_combine(const u64* __restrict__ acc, const u64* __restrict__ conv, u64* __restrict__ out,
                                                         const B200Mod* mods, const ulonglong2* pinv, u32 N, u32 L) {
    const u32 limb = blockIdx.y;
    const u64 q = mods[limb].q;
    const ulonglong2 w = pinv[limb];
    const size_t i2 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i2 * 2 >= N) return;
    const size_t off = (((size_t)blockIdx.z * L + limb) * N) / 2 + i2;
    const ulonglong2 a = __ldcs(reinterpret_cast<const ulonglong2*>(acc) + off);
    const ulonglong2 b = __ldcs(reinterpret_cast<const ulonglong2*>(conv) + off);
    const u64 nq = 0 - q;
    // inputs are canonical (pmult and the forward NTT both end in [0,q))
    const u64 dx = a.x - b.x + (a.x < b.x ? q : 0), dy = a.y - b.y + (a.y < b.y ? q : 0);
    ulonglong2 r;
    r.x = canon2(shoup_mul(dx, w.x, w.y, nq), q);
    r.y = canon2(shoup_mul(dy, w.x, w.y, nq), q);
    reinterpret_cast<ulonglong2*>(out)[off] = r;
}

}  // namespace b200

extern "C" int b200fhe_keyswitch_destroy(b200fhe_keyswitch* k) {
    if (!k) return 0;
    if (k->ctx) cudaSetDevice(k->ctx->device);
    for (b200fhe_bconv* p : k->up) b200fhe_bconv_destroy(p);
    b200fhe_bconv_destroy(k->down);
    if (k->d_pinv) cudaFree(k->d_pinv);
    if (k->d_mods_scaled) cudaFree(k->d_mods_scaled);
    if (k->d_up_map) cudaFree(k->d_up_map);
    if (k->d_ws) cudaFree(k->d_ws);
    delete k;
    return 0;
}

extern "C" int b200fhe_keyswitch_create(b200fhe_ctx* c, int n_q, int n_p, int dnum, b200fhe_keyswitch** out) {
    return b200fhe_keyswitch_create_bgv(c, n_q, n_p, dnum, 1, out);
}

// plain_modulus t > 1: BGV flavour -- ModDown subtracts t [acc t^-1]_P instead of [acc]_P, so the quotient differs from
// acc / P by a multiple of t only and the plaintext survives (switching keys then carry t e_d as their noise term).
extern "C" int b200fhe_keyswitch_create_bgv(b200fhe_ctx* c, int n_q, int n_p, int dnum, uint64_t plain_modulus, b200fhe_keyswitch** out) {
    if (plain_modulus < 1) return b200_set_err(B200FHE_EINVAL, "keyswitch_create: plain modulus must be >= 1");
    if (!c || !out || n_q < 1 || n_p < 1 || dnum < 0) return b200_set_err(B200FHE_EINVAL, "keyswitch_create: bad arguments");
    if (n_q + n_p != c->L) return b200_set_err(B200FHE_EINVAL, "keyswitch_create: context holds %d moduli, need n_q + n_p = %d", c->L, n_q + n_p);
    if (dnum && n_q % dnum) return b200_set_err(B200FHE_EINVAL, "keyswitch_create: n_q=%d not divisible by dnum=%d", n_q, dnum);
    if (c->N < 2) return b200_set_err(B200FHE_EINVAL, "keyswitch_create: N too small");
    for (int i = 0; i < c->L; i++)
        if (!c->moduli[i] || !c->has_inverse[i]) return b200_set_err(B200FHE_EINVAL, "keyswitch_create: modulus %d has no inverse tables", i);
    const int alpha = dnum ? n_q / dnum : 0, total = n_q + n_p;      // dnum == 0: ModDown-only plan (rescale / mod-switch)
    if (alpha > 64 || n_p > 64) return b200_set_err(B200FHE_EUNSUPPORTED, "keyswitch_create: digit or special base larger than 64 limbs");
    b200fhe_keyswitch* k = new b200fhe_keyswitch();
    k->ctx = c; k->n_q = n_q; k->n_p = n_p; k->dnum = dnum; k->alpha = alpha;
    int rc = 0;
    for (int d = 0; d < dnum && !rc; d++) {
        std::vector<int> src, dst;
        for (int g = d * alpha; g < (d + 1) * alpha; g++) src.push_back(g);
        for (int g = 0; g < total; g++)
            if (g < d * alpha || g >= (d + 1) * alpha) dst.push_back(g);
        b200fhe_bconv* p = nullptr;
        rc = b200fhe_bconv_create(c, src.data(), alpha, dst.data(), (int)dst.size(), &p);
        if (!rc) k->up.push_back(p);
    }
    if (!rc) {
        std::vector<int> src, dst;
        for (int g = n_q; g < total; g++) src.push_back(g);
        for (int g = 0; g < n_q; g++) dst.push_back(g);
        rc = b200fhe_bconv_create_scaled(c, src.data(), n_p, dst.data(), n_q, plain_modulus, &k->down);
    }
    if (rc) {
        b200fhe_keyswitch_destroy(k);
        return rc;
    }
    std::vector<ulonglong2> pinv(n_q);
    for (int l = 0; l < n_q; l++) {
        const uint64_t q = c->moduli[l];
        uint64_t P = 1;
        for (int j = n_q; j < total; j++) P = nt::mulmod(P, c->moduli[j] % q, q);
        const uint64_t iv = nt::invmod(P, q);
        pinv[l] = make_ulonglong2(iv, nt::shoup(iv, q));
    }
    // scale step of every conversion folded into the inverse NTT that precedes it: limb l of Q belongs to exactly one digit,
    // the limbs of P to the ModDown conversion
    std::vector<B200Mod> ms(c->h_mods.begin(), c->h_mods.begin() + total);
    for (int l = 0; l < total; l++) {
        const uint64_t q = c->moduli[l];
        int lo, hi;                                                   // the source base this limb is converted from
        uint64_t extra = 1;
        if (l < n_q) {
            if (!dnum) continue;
            lo = (l / alpha) * alpha; hi = lo + alpha;
        } else {
            lo = n_q; hi = total; extra = plain_modulus % q;
        }
        uint64_t h = extra;
        for (int i = lo; i < hi; i++)
            if (i != l) h = nt::mulmod(h, c->moduli[i] % q, q);
        const uint64_t iv = nt::invmod(h, q);
        ms[l].ninv = nt::mulmod(ms[l].ninv, iv, q);
        ms[l].ninv_s = nt::shoup(ms[l].ninv, q);
        ms[l].iw1n = nt::mulmod(ms[l].iw1n, iv, q);
        ms[l].iw1n_s = nt::shoup(ms[l].iw1n, q);
    }
    const size_t N = c->N;
    size_t o = 0;
    k->o_x = o;    o += dnum ? (size_t)n_q * N : 0;                                  // INTT(c)
    k->o_y = o;    o += (size_t)(dnum && n_q > 2 * n_p ? n_q : 2 * n_p) * N;         // scaled digits; reused for the scaled P part of ModDown
    k->o_up = o;   o += (size_t)dnum * total * N;                                    // raised digits
    k->o_accq = o; o += dnum ? (size_t)2 * n_q * N : 0;
    k->o_accp = o; o += dnum ? (size_t)2 * n_p * N : 0;
    k->o_conv = o; o += (size_t)2 * n_q * N;
    k->ws_words = o;
    cudaError_t e = cudaSetDevice(c->device);
    if (e == cudaSuccess) e = cudaMalloc(&k->d_pinv, sizeof(ulonglong2) * n_q);
    if (e == cudaSuccess) e = cudaMalloc(&k->d_ws, sizeof(uint64_t) * o);
    if (e == cudaSuccess) e = cudaMalloc(&k->d_mods_scaled, sizeof(B200Mod) * total);
    if (e == cudaSuccess) e = cudaMemcpy(k->d_mods_scaled, ms.data(), sizeof(B200Mod) * total, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(k->d_pinv, pinv.data(), sizeof(ulonglong2) * n_q, cudaMemcpyHostToDevice);
    if (dnum) {
        const int alpha = n_q / dnum;
        for (int d = 0; d < dnum; d++)
            for (int g = 0; g < total; g++)
                if (g < d * alpha || g >= (d + 1) * alpha) k->up_map.push_back(make_uint2((unsigned)(d * total + g), (unsigned)g));
        if (e == cudaSuccess) e = cudaMalloc(&k->d_up_map, sizeof(uint2) * k->up_map.size());
        if (e == cudaSuccess) e = cudaMemcpy(k->d_up_map, k->up_map.data(), sizeof(uint2) * k->up_map.size(), cudaMemcpyHostToDevice);
    }
    if (e != cudaSuccess) {
        b200fhe_keyswitch_destroy(k);
        return b200_set_err(e == cudaErrorMemoryAllocation ? B200FHE_ENOMEM : B200FHE_ECUDA, "keyswitch_create: %s", cudaGetErrorString(e));
    }
    *out = k;
    return 0;
}

// steps 1-2: every digit of c raised to the full basis, NTT domain.  d_up [dnum][n_q + n_p][N].
int b200_bconv_modup_protected_split(b200fhe_bconv* p, const uint64_t* d_in, uint64_t* d_y, uint64_t* d_out, uint64_t* d_out_hi, size_t k_split,
                                     uint32_t* d_verdict, uint64_t* d_sums, const b200fhe_bconv_inject* inject, size_t n_inject, void* stream);

// verdict (optional): [dnum][n_q + n_p] checksum verdicts of the digits' conversions (b200fhe_bconv_modup_protected layout:
// alpha source-limb entries, then the targets)
// vntt (optional, only with verdict): per-limb verdicts of the protected transforms, in launch order: INTT(c) [n_q], then per digit
// its targets below [d alpha] and above [total - (d + 1) alpha]
// skip_own: leave the digits' own rows of d_up unwritten (apply reads those limbs from c itself)
static int modup_impl(b200fhe_keyswitch* k, const uint64_t* d_c, uint64_t* d_up, uint32_t* verdict, void* stream, uint32_t* vntt = nullptr,
                      bool skip_own = false) {
    if (!k || !d_c || !d_up) return b200_set_err(B200FHE_EINVAL, "keyswitch_modup: null argument");
    if (!k->dnum) return b200_set_err(B200FHE_EINVAL, "keyswitch_modup: this is a ModDown-only plan (dnum = 0)");
    b200fhe_ctx* c = k->ctx;
    CK(cudaSetDevice(c->device));
    cudaStream_t s = (cudaStream_t)stream;
    const size_t N = c->N, a = (size_t)k->alpha, total = (size_t)k->n_q + k->n_p;
    uint64_t* x = k->d_ws + k->o_x;
    uint64_t* y = k->d_ws + k->o_y;
    // x = INTT(c): out of place where the kernels allow it (plain switch, two-pass sizes), else a device copy and the in-place form
    bool x_done = false;
    if (!verdict) {
        const int rc = b200_nwt_backward_mods_oop(c, d_c, x, 1, (size_t)k->n_q, 0, k->d_mods_scaled, stream);
        if (rc < 0) return rc;
        x_done = rc == 0;
    }
    if (!x_done) CK(cudaMemcpyAsync(x, d_c, sizeof(uint64_t) * k->n_q * N, cudaMemcpyDeviceToDevice, s));
    // the digit's own limbs pass through
    for (int d = 0; d < k->dnum && !skip_own; d++)
        CK(cudaMemcpyAsync(d_up + ((size_t)d * total + (size_t)d * a) * N, d_c + (size_t)d * a * N, sizeof(uint64_t) * a * N,
                           cudaMemcpyDeviceToDevice, s));
    // unprotected: the inverse NTT's final constant carries the conversion's scale step, so x already IS y = x (D/q_j)^-1;
    // protected: plain inverse NTT, the scale step runs as its own checked kernel
    if (verdict && vntt) { RC(b200fhe_nwt_backward_protected(c, x, 1, (size_t)k->n_q, 0, vntt, nullptr, stream)); vntt += k->n_q; }
    else if (verdict) RC(b200fhe_nwt_backward(c, x, 1, (size_t)k->n_q, 0, stream));
    else if (!x_done) RC(b200_nwt_backward_mods(c, x, 1, (size_t)k->n_q, 0, k->d_mods_scaled, stream));
    // the forward transforms of all digits' converted limbs go out as one scatter launch (not for the checked transforms, whose
    // verdicts are per range, nor for the one-kernel sizes)
    static const bool no_merge = getenv("B200FHE_KS_NO_MERGE") != nullptr;
    const bool merged = !vntt && !no_merge && c->log_n >= 12 && k->up_map.size() <= 65535 && k->d_up_map;
    for (int d = 0; d < k->dnum; d++) {
        b200fhe_bconv* p = k->up[d];
        const uint64_t* xd = x + (size_t)d * a * N;
        uint64_t* yd = y + (size_t)d * a * N;
        uint64_t* upd = (uint64_t*)d_up + (size_t)d * total * N;
        const size_t below = (size_t)d * a, above = total - below - a;
        if (verdict) {
            const bool hit = k->inj_conv == d;
            RC(b200_bconv_modup_protected_split(p, xd, yd, upd, upd + (below + a) * N, below, verdict + (size_t)d * total, nullptr,
                                                hit ? k->inj.data() : nullptr, hit ? k->inj.size() : 0, stream));
        } else {
            // one launch for the targets below and above the digit (its own limbs are the gap); two range calls where the
            // tcgen05 contraction declines the shape
            const int one = b200_bconv_contract_skip(p, xd, upd, below, a, stream);
            if (one < 0) return one;
            if (one > 0) {
                RC(b200fhe_bconv_contract(p, xd, upd, 0, below, stream));
                RC(b200fhe_bconv_contract(p, xd, upd + (below + a) * N, below, above, stream));
            }
        }
        if (vntt) {
            if (below) RC(b200fhe_nwt_forward_protected(c, upd, 1, below, 0, vntt, nullptr, nullptr, 0, stream));
            RC(b200fhe_nwt_forward_protected(c, upd + (below + a) * N, 1, above, below + a, vntt + below, nullptr, nullptr, 0, stream));
            vntt += below + above;
        } else if (!merged) {
            if (below) RC(b200fhe_nwt_forward(c, upd, 1, below, 0, stream));
            RC(b200fhe_nwt_forward(c, upd + (below + a) * N, 1, above, below + a, stream));
        }
    }
    if (merged) {
        const int rc = b200_nwt_forward_map(c, d_up, k->d_up_map, k->up_map.data(), k->up_map.size(), stream);
        if (rc < 0) return rc;
        if (rc > 0) return b200_set_err(B200FHE_EUNSUPPORTED, "keyswitch_modup: scatter transform unavailable");
    }
    return 0;
}

extern "C" int b200fhe_keyswitch_modup(b200fhe_keyswitch* k, const uint64_t* d_c, uint64_t* d_up, void* stream) {
    return modup_impl(k, d_c, d_up, nullptr, stream);
}

// step 4 for n_poly polynomials: d_acc_q [n_poly][n_q][N], d_acc_p [n_poly][n_p][N] (overwritten), d_out [n_poly][n_q][N]
// (may alias d_acc_q).  n_poly <= 2.
// convert_only: stop after conv = NTT(Conv_{P->Q}(INTT(acc_p))) (left in the workspace); the caller fuses the last step elsewhere
static int moddown_impl(b200fhe_keyswitch* k, const uint64_t* d_acc_q, uint64_t* d_acc_p, uint64_t* d_out, size_t n_poly,
                        uint32_t* verdict /*optional [n_poly][n_p + n_q]*/, void* stream, uint32_t* vntt = nullptr /*[n_poly n_p] then [n_poly n_q]*/,
                        bool convert_only = false) {
    if (!k || !d_acc_q || !d_acc_p || !d_out) return b200_set_err(B200FHE_EINVAL, "keyswitch_moddown: null argument");
    if (n_poly < 1 || n_poly > 2) return b200_set_err(B200FHE_EINVAL, "keyswitch_moddown: n_poly must be 1 or 2");
    b200fhe_ctx* c = k->ctx;
    CK(cudaSetDevice(c->device));
    cudaStream_t s = (cudaStream_t)stream;
    const size_t N = c->N, nq = (size_t)k->n_q, np = (size_t)k->n_p;
    uint64_t* y = k->d_ws + k->o_y;
    uint64_t* conv = k->d_ws + k->o_conv;
    if (verdict && vntt) RC(b200fhe_nwt_backward_protected(c, d_acc_p, n_poly, np, nq, vntt, nullptr, stream));
    else if (verdict) RC(b200fhe_nwt_backward(c, d_acc_p, n_poly, np, nq, stream));
    else RC(b200_nwt_backward_mods(c, d_acc_p, n_poly, np, nq, k->d_mods_scaled, stream));     // scale step folded in: acc_p is y
    for (size_t p = 0; p < n_poly; p++) {
        if (verdict) {
            const bool hit = k->inj_conv == k->dnum + (int)p;
            RC(b200_bconv_modup_protected_split(k->down, d_acc_p + p * np * N, y + p * np * N, conv + p * nq * N, nullptr, 0,
                                                verdict + p * (np + nq), nullptr, hit ? k->inj.data() : nullptr, hit ? k->inj.size() : 0, stream));
        } else {
            RC(b200fhe_bconv_contract(k->down, d_acc_p + p * np * N, conv + p * nq * N, 0, nq, stream));
        }
    }
    if (verdict && vntt) RC(b200fhe_nwt_forward_protected(c, conv, n_poly, nq, 0, vntt + n_poly * np, nullptr, nullptr, 0, stream));
    else RC(b200fhe_nwt_forward(c, conv, n_poly, nq, 0, stream));
    if (convert_only) return 0;
    dim3 grid((unsigned)((N / 2 + 255) / 256), (unsigned)nq, (unsigned)n_poly);
    k_moddown_combine<<<grid, 256, 0, s>>>((const u64*)d_acc_q, (const u64*)conv, (u64*)d_out, c->d_mods, k->d_pinv, (u32)N, (u32)nq);
    b200_count_launches(1);
    CK(cudaGetLastError());
    return 0;
}

extern "C" int b200fhe_keyswitch_moddown(b200fhe_keyswitch* k, const uint64_t* d_acc_q, uint64_t* d_acc_p, uint64_t* d_out,
                                         size_t n_poly, void* stream) {
    return moddown_impl(k, d_acc_q, d_acc_p, d_out, n_poly, nullptr, stream);
}

// The whole switch: d_c [n_q][N]; d_evk_q [dnum][2][n_q][N]; d_evk_p [dnum][2][n_p][N]; d_out [2][n_q][N].
// out[0] + out[1] s  ~  c s'   when evk_d = (-a_d s + e_d + P B_d s', a_d) with B_d the CRT basis element of digit d.
static size_t modup_ntt_count(const b200fhe_keyswitch* k) { return (size_t)k->n_q + (size_t)k->dnum * ((size_t)k->n_q + k->n_p - k->alpha); }
static int apply_impl(b200fhe_keyswitch* k, const uint64_t* d_c, const uint64_t* d_evk_q, const uint64_t* d_evk_p, uint64_t* d_out,
                      uint32_t* verdict, void* stream, uint32_t* vntt = nullptr) {
    if (!k || !d_c || !d_evk_q || !d_evk_p || !d_out) return b200_set_err(B200FHE_EINVAL, "keyswitch_apply: null argument");
    if (!k->dnum) return b200_set_err(B200FHE_EINVAL, "keyswitch_apply: this is a ModDown-only plan (dnum = 0)");
    b200fhe_ctx* c = k->ctx;
    const size_t N = c->N, nq = (size_t)k->n_q, np = (size_t)k->n_p, total = nq + np;
    uint64_t* up = k->d_ws + k->o_up;
    uint64_t* accq = k->d_ws + k->o_accq;
    uint64_t* accp = k->d_ws + k->o_accp;
    static const bool no_fuse = getenv("B200FHE_KS_NO_FUSED_COMBINE") != nullptr;
    const bool fused = k->dnum <= 4 && !no_fuse;
    RC(modup_impl(k, d_c, up, verdict, stream, vntt, fused));
    std::vector<const uint64_t*> ct(k->dnum), pt(k->dnum);
    for (int d = 0; d < k->dnum; d++) { ct[d] = d_evk_q + (size_t)d * 2 * nq * N; pt[d] = up + (size_t)d * total * N; }
    if (fused) {
        // the P part first: its ModDown conversion only needs acc_p; then the Q part's inner product with ModDown's last step
        // (subtract, times P^-1) as its epilogue, straight into d_out -- acc_q never exists in memory
        for (int d = 0; d < k->dnum; d++) { ct[d] = d_evk_p + (size_t)d * 2 * np * N; pt[d] = up + ((size_t)d * total + nq) * N; }
        RC(b200fhe_pmult_accum(c, ct.data(), pt.data(), (size_t)k->dnum, accp, 2, np, nq, 0, stream));
        RC(moddown_impl(k, accq, accp, d_out, 2, verdict ? verdict + (size_t)k->dnum * total : nullptr, stream,
                        vntt ? vntt + modup_ntt_count(k) : nullptr, true));
        for (int d = 0; d < k->dnum; d++) { ct[d] = d_evk_q + (size_t)d * 2 * nq * N; pt[d] = up + (size_t)d * total * N; }
        if (launch_pmult_moddown((const u64* const*)ct.data(), (const u64* const*)pt.data(), (size_t)k->dnum, (const u64*)(k->d_ws + k->o_conv),
                                 (u64*)d_out, c->d_mods, k->d_pinv, (u32)N, (u32)nq, (cudaStream_t)stream, (const u64*)d_c, (u32)k->alpha))
            return b200_set_err(B200FHE_EUNSUPPORTED, "keyswitch_apply: fused inner product unavailable");
        b200_count_launches(1);
        CK(cudaGetLastError());
        return 0;
    }
    RC(b200fhe_pmult_accum(c, ct.data(), pt.data(), (size_t)k->dnum, accq, 2, nq, 0, 0, stream));
    for (int d = 0; d < k->dnum; d++) { ct[d] = d_evk_p + (size_t)d * 2 * np * N; pt[d] = up + ((size_t)d * total + nq) * N; }
    RC(b200fhe_pmult_accum(c, ct.data(), pt.data(), (size_t)k->dnum, accp, 2, np, nq, 0, stream));
    return moddown_impl(k, accq, accp, d_out, 2, verdict ? verdict + (size_t)k->dnum * total : nullptr, stream, vntt ? vntt + modup_ntt_count(k) : nullptr);
}

extern "C" int b200fhe_keyswitch_apply(b200fhe_keyswitch* k, const uint64_t* d_c, const uint64_t* d_evk_q, const uint64_t* d_evk_p,
                                       uint64_t* d_out, void* stream) {
    return apply_impl(k, d_c, d_evk_q, d_evk_p, d_out, nullptr, stream);
}

// Same switch with every base conversion checksum-carrying (b200fhe_bconv_modup_protected: scale, memory and contraction
// checks).  d_verdict [(dnum + 2)][n_q + n_p]: rows 0..dnum-1 = the digits' ModUp conversions (alpha source entries, then
// the targets in ascending limb order), rows dnum, dnum+1 = the ModDown conversions of out[0], out[1] (n_p source entries,
// then n_q targets).  All zero = no check failed.  Needs digits and special base of at most 16 limbs.
extern "C" int b200fhe_keyswitch_apply_protected(b200fhe_keyswitch* k, const uint64_t* d_c, const uint64_t* d_evk_q, const uint64_t* d_evk_p,
                                                 uint64_t* d_out, uint32_t* d_verdict, void* stream) {
    if (!d_verdict) return b200_set_err(B200FHE_EINVAL, "keyswitch_apply_protected: null verdict");
    int rc = apply_impl(k, d_c, d_evk_q, d_evk_p, d_out, d_verdict, stream);
    if (k) { k->inj_conv = -1; k->inj.clear(); }              // an armed injection is consumed by one call
    return rc;
}

// The same, with EVERY limb-transform of the switch protected as well (b200fhe_nwt_forward_protected / _backward_protected): the
// 320 transforms at C5's shape are 48 % of a switch's cost by the reference's own split (profile_framewk/draw_dnum.py:5-9).
// d_verdict_ntt [b200fhe_keyswitch_ntt_count]: INTT(c) [n_q]; per digit its targets below then above the digit; ModDown's INTT of the
// P part [2 n_p] and NTT of the converted part [2 n_q].  All zero = no transform's checksum tripped.
extern "C" int b200fhe_keyswitch_apply_protected_full(b200fhe_keyswitch* k, const uint64_t* d_c, const uint64_t* d_evk_q, const uint64_t* d_evk_p,
                                                      uint64_t* d_out, uint32_t* d_verdict, uint32_t* d_verdict_ntt, void* stream) {
    if (!d_verdict || !d_verdict_ntt) return b200_set_err(B200FHE_EINVAL, "keyswitch_apply_protected_full: null verdict");
    int rc = apply_impl(k, d_c, d_evk_q, d_evk_p, d_out, d_verdict, stream, d_verdict_ntt);
    if (k) { k->inj_conv = -1; k->inj.clear(); }
    return rc;
}
extern "C" int b200fhe_keyswitch_ntt_count(const b200fhe_keyswitch* k, size_t* n) {
    if (!k || !n || !k->dnum) return b200_set_err(B200FHE_EINVAL, "keyswitch_ntt_count: needs a full key-switch plan");
    *n = modup_ntt_count(k) + 2 * (size_t)k->n_p + 2 * (size_t)k->n_q;
    return 0;
}

// Arms faults for the NEXT b200fhe_keyswitch_apply_protected: `conversion` in [0, dnum) = that digit's ModUp conversion,
// dnum / dnum + 1 = the ModDown conversion of out[0] / out[1]; entries as in b200fhe_bconv_modup_protected.
extern "C" int b200fhe_keyswitch_set_inject(b200fhe_keyswitch* k, int conversion, const b200fhe_bconv_inject* inject, size_t n_inject) {
    if (!k || conversion < 0 || conversion >= k->dnum + 2 || n_inject > B200FHE_BCONV_MAX_INJECT || (n_inject && !inject))
        return b200_set_err(B200FHE_EINVAL, "keyswitch_set_inject: bad arguments");
    k->inj_conv = n_inject ? conversion : -1;
    k->inj.assign(inject, inject + n_inject);
    return 0;
}

extern "C" int b200fhe_keyswitch_info(const b200fhe_keyswitch* k, int* n_q, int* n_p, int* dnum, size_t* workspace_bytes) {
    if (!k) return b200_set_err(B200FHE_EINVAL, "keyswitch_info: null plan");
    if (n_q) *n_q = k->n_q;
    if (n_p) *n_p = k->n_p;
    if (dnum) *dnum = k->dnum;
    if (workspace_bytes) *workspace_bytes = k->ws_words * sizeof(uint64_t);
    return 0;
}
