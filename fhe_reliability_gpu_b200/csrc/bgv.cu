// BGV host roles and evaluator composition behind the C ABI: what reliability_test/dotprod_test.cu:67-185
// obtains from Phantom's PhantomContext / PhantomSecretKey / PhantomPublicKey / PhantomBatchEncoder and the free functions
// multiply, relinearize_inplace, mod_switch_to_next_inplace, rotate_inplace, add_inplace (symbol list: SURVEY 8(b); Phantom's
// sources are absent from the reference tree, so this follows the published BGV / SEAL conventions the harness relies on:
// BatchEncoder slot order with generator 3, rotate by step = Galois element 3^step, hybrid key-switch, the correction factor
// of mod_switch_to_next).  shim/phantom.h wraps these entries in the Phantom class names so dotprod_test.cu compiles unmodified.
//
// Every polynomial operation runs on the GPU through this library's own kernels (NTT, ct x pt multiply-accumulate, tensor,
// Galois permutation, key-switch, ModDown); the host only samples (secret, noise, uniform masks), encodes / decodes the slot
// order, and recombines the decrypted residues (Garner) -- there is no CPU arithmetic fallback.
//
// Levels: a ciphertext with l data limbs lives under the moduli q_0..q_{l-1}; its key-switches use the context
// (q_0..q_{l-1}, p_0..p_{n_p-1}) with dnum(l) = the largest divisor count such that a digit has at most n_p limbs.  Switching
// keys are made on first use per (source secret, level) and cached in the object, which holds the secret key (this is a
// fault-injection harness, not a deployment: PhantomRelinKey / PhantomGaloisKey in the shim are handles into this cache).
#include "../../include/b200fhe.h"
#include "ctx.hpp"
#include "numth.hpp"
#include <cmath>
#include <map>
#include <random>
#include <vector>

int b200_set_err(int code, const char* fmt, ...);
namespace nt = b200::nt;

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

namespace {
typedef unsigned __int128 u128;
constexpr uint32_t kGen = 3;                     // SEAL / Phantom GaloisTool generator

struct EvKey { uint64_t* q = nullptr; uint64_t* p = nullptr; };      // [dnum][2][l][N], [dnum][2][n_p][N]

struct Level {                                    // everything tied to "l data limbs"
    int l = 0, dnum = 0;
    b200fhe_ctx* ctx = nullptr;                   // q_0..q_{l-1}, p_0..p_{n_p-1}
    b200fhe_keyswitch* ks = nullptr;
    b200fhe_ctx* qctx = nullptr;                  // q_0..q_{l-1} only (mod_switch: the last prime plays the special prime)
    b200fhe_keyswitch* down = nullptr;            // ModDown-only plan (l - 1, 1)
    uint64_t* d_s = nullptr;                      // secret key, NTT form, [l + n_p][N]
    uint64_t* d_ones = nullptr;                   // the constant polynomial 1 in NTT form, [l + n_p][N]
    std::map<uint32_t, EvKey> keys;               // source secret: 0 = s^2 (relinearisation), g = s(x^g)
};
}  // namespace

struct b200fhe_bgv {
    int log_n = 0, n_q = 0, n_p = 0, device = 0;
    size_t N = 0;
    uint64_t t = 0;
    std::vector<uint64_t> moduli;                 // n_q data primes, then n_p special primes
    b200fhe_ctx* tctx = nullptr;                  // the plain modulus as an NTT modulus (batching)
    std::vector<uint32_t> index_map;              // BatchEncoder matrix_reps_index_map
    std::vector<int8_t> s;                        // ternary secret
    bool have_sk = false, have_pk = false;
    uint64_t* d_pk = nullptr;                     // [2][n_q][N]
    std::vector<Level> lv;                        // index l (1..n_q)
    std::mt19937_64 rng;
    uint64_t* d_tmp = nullptr;                    // scratch, 8 x (n_q + n_p) x N words
    std::vector<uint64_t> h_tmp;
};

namespace {

int level_dnum(int l, int n_p) {
    for (int alpha = n_p < l ? n_p : l; alpha >= 1; alpha--)
        if (l % alpha == 0) return l / alpha;
    return l;
}

void free_level(Level& L) {
    for (auto& kv : L.keys) { if (kv.second.q) cudaFree(kv.second.q); if (kv.second.p) cudaFree(kv.second.p); }
    L.keys.clear();
    if (L.ks) b200fhe_keyswitch_destroy(L.ks);
    if (L.down) b200fhe_keyswitch_destroy(L.down);
    if (L.ctx) b200fhe_ctx_destroy(L.ctx);
    if (L.qctx) b200fhe_ctx_destroy(L.qctx);
    if (L.d_s) cudaFree(L.d_s);
    if (L.d_ones) cudaFree(L.d_ones);
    L = Level();
}

// global index of limb i of level l's key-switch context
inline int level_limb(const b200fhe_bgv* h, int l, int i) { return i < l ? i : h->n_q + (i - l); }

// small signed coefficients -> residues on the level context's limbs [first, first + count), uploaded and transformed
int upload_signed_ntt(b200fhe_bgv* h, Level& L, const std::vector<int64_t>& v, int first, int count, uint64_t* d_out) {
    const size_t N = h->N;
    h->h_tmp.resize((size_t)count * N);
    for (int i = 0; i < count; i++) {
        const uint64_t q = h->moduli[level_limb(h, L.l, first + i)];
        for (size_t k = 0; k < N; k++) {
            const int64_t x = v[k];
            h->h_tmp[(size_t)i * N + k] = x >= 0 ? (uint64_t)x % q : q - ((uint64_t)(-x) % q);
            if (h->h_tmp[(size_t)i * N + k] == q) h->h_tmp[(size_t)i * N + k] = 0;
        }
    }
    CK(cudaMemcpy(d_out, h->h_tmp.data(), sizeof(uint64_t) * count * N, cudaMemcpyHostToDevice));
    return b200fhe_nwt_forward(L.ctx, d_out, 1, (size_t)count, (size_t)first, nullptr);
}

// out[i] = sum_k a_k[i] (.) b_k[i] over `count` limbs starting at `first` of the level context (all NTT form)
int mac(Level& L, std::initializer_list<std::pair<const uint64_t*, const uint64_t*>> terms, uint64_t* d_out, int first, int count) {
    std::vector<const uint64_t*> a, b;
    for (auto& t : terms) { a.push_back(t.first); b.push_back(t.second); }
    return b200fhe_pmult_accum(L.ctx, a.data(), b.data(), a.size(), d_out, 1, (size_t)count, (size_t)first, 0, nullptr);
}

std::vector<int64_t> sample_noise(b200fhe_bgv* h, int64_t scale) {      // discrete Gaussian sigma 3.2, clipped at 6 sigma, times `scale`
    std::normal_distribution<double> nd(0.0, 3.2);
    std::vector<int64_t> e(h->N);
    for (auto& x : e) {
        double v;
        do { v = nd(h->rng); } while (std::fabs(v) > 19.2);
        x = (int64_t)std::llround(v) * scale;
    }
    return e;
}
std::vector<int64_t> sample_ternary(b200fhe_bgv* h) {
    std::vector<int64_t> u(h->N);
    for (auto& x : u) x = (int64_t)(h->rng() % 3) - 1;
    return u;
}
uint64_t uniform_mod(std::mt19937_64& g, uint64_t q) {                  // rejection sampling
    const uint64_t lim = ~0ull - (~0ull % q + 1) % q;
    uint64_t x;
    do { x = g(); } while (x > lim);
    return x % q;
}

}  // namespace

// ---- levels -------------------------------------------------------------------------------------------------------
static int ensure_level(b200fhe_bgv* h, int l, Level** out) {
    if (l < 1 || l > h->n_q) return b200_set_err(B200FHE_EINVAL, "bgv: a ciphertext must have 1..%d limbs, got %d", h->n_q, l);
    Level& L = h->lv[l];
    if (!L.ctx) {
        CK(cudaSetDevice(h->device));
        const int total = l + h->n_p;
        std::vector<uint64_t> m(total);
        for (int i = 0; i < total; i++) m[i] = h->moduli[level_limb(h, l, i)];
        L.l = l; L.dnum = level_dnum(l, h->n_p);
        RC(b200fhe_ctx_create(h->log_n, m.data(), total, h->device, &L.ctx));
        RC(b200fhe_keyswitch_create_bgv(L.ctx, l, h->n_p, L.dnum, h->t, &L.ks));
        if (l >= 2) {
            RC(b200fhe_ctx_create(h->log_n, m.data(), l, h->device, &L.qctx));
            RC(b200fhe_keyswitch_create_bgv(L.qctx, l - 1, 1, 0, h->t, &L.down));
        }
        CK(cudaMalloc(&L.d_ones, sizeof(uint64_t) * total * h->N));
        std::vector<uint64_t> ones((size_t)total * h->N, 1);
        CK(cudaMemcpy(L.d_ones, ones.data(), sizeof(uint64_t) * ones.size(), cudaMemcpyHostToDevice));
    }
    if (!L.d_s && h->have_sk) {
        CK(cudaMalloc(&L.d_s, sizeof(uint64_t) * (l + h->n_p) * h->N));
        std::vector<int64_t> sv(h->s.begin(), h->s.end());
        RC(upload_signed_ntt(h, L, sv, 0, l + h->n_p, L.d_s));
    }
    *out = &L;
    return 0;
}

extern "C" int b200fhe_bgv_destroy(b200fhe_bgv* h) {
    if (!h) return 0;
    cudaSetDevice(h->device);
    for (auto& L : h->lv) free_level(L);
    if (h->tctx) b200fhe_ctx_destroy(h->tctx);
    if (h->d_pk) cudaFree(h->d_pk);
    if (h->d_tmp) cudaFree(h->d_tmp);
    delete h;
    return 0;
}

// coeff_moduli: n_moduli primes of which the LAST special_size are the key-switching (special) primes, as
// EncryptionParameters::set_coeff_modulus + set_special_modulus_size give them (dotprod_test.cu:201-205);
// plain_modulus: a prime = 1 mod 2N (PlainModulus::Batching, :206).
extern "C" int b200fhe_bgv_create(int log_n, const uint64_t* coeff_moduli, int n_moduli, int special_size, uint64_t plain_modulus,
                                  uint64_t seed, int device, b200fhe_bgv** out) {
    if (!coeff_moduli || !out || special_size < 1 || n_moduli <= special_size || log_n < 1 || log_n > 17)
        return b200_set_err(B200FHE_EINVAL, "bgv_create: bad arguments");
    const uint64_t N = 1ull << log_n;
    if (plain_modulus < 3 || !nt::is_prime(plain_modulus) || (plain_modulus - 1) % (2 * N))
        return b200_set_err(B200FHE_EINVAL, "bgv_create: plain modulus %llu is not a batching prime for N=%llu", (unsigned long long)plain_modulus,
                            (unsigned long long)N);
    b200fhe_bgv* h = new b200fhe_bgv();
    h->log_n = log_n; h->N = N; h->n_p = special_size; h->n_q = n_moduli - special_size; h->t = plain_modulus; h->device = device;
    h->moduli.assign(coeff_moduli, coeff_moduli + n_moduli);
    h->lv.resize(h->n_q + 1);
    h->rng.seed(seed);
    int rc = b200fhe_ctx_create(log_n, &plain_modulus, 1, device, &h->tctx);
    cudaError_t e = rc ? cudaSuccess : cudaMalloc(&h->d_tmp, sizeof(uint64_t) * 8 * n_moduli * N);
    if (rc || e != cudaSuccess) {
        b200fhe_bgv_destroy(h);
        return rc ? rc : b200_set_err(B200FHE_ENOMEM, "bgv_create: %s", cudaGetErrorString(e));
    }
    // BatchEncoder::matrix_reps_index_map: slot i of row 0 sits at the evaluation point 3^i, row 1 at -3^i, bit-reversed order
    h->index_map.resize(N);
    uint64_t pos = 1;
    for (uint64_t i = 0; i < N / 2; i++) {
        h->index_map[i] = (uint32_t)nt::bitrev((pos - 1) >> 1, log_n);
        h->index_map[N / 2 + i] = (uint32_t)nt::bitrev((2 * N - pos - 1) >> 1, log_n);
        pos = pos * kGen % (2 * N);
    }
    *out = h;
    return 0;
}

extern "C" int b200fhe_bgv_info(const b200fhe_bgv* h, int* log_n, int* n_q, int* n_p, uint64_t* plain_modulus, uint64_t* moduli) {
    if (!h) return b200_set_err(B200FHE_EINVAL, "bgv_info: null");
    if (log_n) *log_n = h->log_n;
    if (n_q) *n_q = h->n_q;
    if (n_p) *n_p = h->n_p;
    if (plain_modulus) *plain_modulus = h->t;
    if (moduli) for (size_t i = 0; i < h->moduli.size(); i++) moduli[i] = h->moduli[i];
    return 0;
}

// ---- key generation -------------------------------------------------------------------------------------------------
// PhantomSecretKey::gen_secretkey (its constructor, dotprod_test.cu:73): ternary secret.
extern "C" int b200fhe_bgv_gen_secretkey(b200fhe_bgv* h) {
    if (!h) return b200_set_err(B200FHE_EINVAL, "bgv_gen_secretkey: null");
    CK(cudaSetDevice(h->device));
    for (auto& L : h->lv) { if (L.d_s) { cudaFree(L.d_s); L.d_s = nullptr; } for (auto& kv : L.keys) { cudaFree(kv.second.q); cudaFree(kv.second.p); } L.keys.clear(); }
    h->s.resize(h->N);
    for (auto& x : h->s) x = (int8_t)((int)(h->rng() % 3) - 1);
    h->have_sk = true; h->have_pk = false;
    return 0;
}

// PhantomSecretKey::gen_publickey (:74): pk = (-(a s) + t e, a) over the data primes.
extern "C" int b200fhe_bgv_gen_publickey(b200fhe_bgv* h) {
    if (!h || !h->have_sk) return b200_set_err(B200FHE_EINVAL, "bgv_gen_publickey: no secret key");
    Level* L;
    RC(ensure_level(h, h->n_q, &L));
    const size_t N = h->N, nq = (size_t)h->n_q;
    if (!h->d_pk) CK(cudaMalloc(&h->d_pk, sizeof(uint64_t) * 2 * nq * N));
    std::vector<uint64_t> a(nq * N), na(nq * N);
    for (size_t i = 0; i < nq; i++)
        for (size_t k = 0; k < N; k++) {
            const uint64_t q = h->moduli[i], x = uniform_mod(h->rng, q);
            a[i * N + k] = x; na[i * N + k] = x ? q - x : 0;
        }
    uint64_t* d_na = h->d_tmp;
    uint64_t* d_e = h->d_tmp + nq * N;
    CK(cudaMemcpy(h->d_pk + nq * N, a.data(), sizeof(uint64_t) * nq * N, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_na, na.data(), sizeof(uint64_t) * nq * N, cudaMemcpyHostToDevice));
    RC(upload_signed_ntt(h, *L, sample_noise(h, (int64_t)h->t), 0, (int)nq, d_e));
    RC(mac(*L, {{d_na, L->d_s}, {d_e, L->d_ones}}, h->d_pk, 0, (int)nq));
    CK(cudaDeviceSynchronize());
    h->have_pk = true;
    return 0;
}

// Switching key from the secret `from` (NTT form on the level's l + n_p limbs) to s, for level L:
// evk_d = (-(a_d s) + t e_d + P B_d from, a_d), B_d = 1 on the limbs of digit d and 0 elsewhere (b200fhe_keyswitch_apply's contract).
static int make_evk(b200fhe_bgv* h, Level& L, const uint64_t* d_from, EvKey* out) {
    const size_t N = h->N;
    const int l = L.l, np = h->n_p, total = l + np, alpha = l / L.dnum;
    EvKey k;
    CK(cudaMalloc(&k.q, sizeof(uint64_t) * L.dnum * 2 * l * N));
    CK(cudaMalloc(&k.p, sizeof(uint64_t) * L.dnum * 2 * np * N));
    uint64_t *d_a = h->d_tmp, *d_na = d_a + (size_t)total * N, *d_e = d_na + (size_t)total * N, *d_f = d_e + (size_t)total * N,
             *d_b = d_f + (size_t)total * N;
    std::vector<uint64_t> a((size_t)total * N), na((size_t)total * N), f((size_t)total * N);
    for (int d = 0; d < L.dnum; d++) {
        for (int i = 0; i < total; i++) {
            const uint64_t q = h->moduli[level_limb(h, l, i)];
            uint64_t P = 0;
            if (i >= d * alpha && i < (d + 1) * alpha) {
                P = 1;
                for (int j = 0; j < np; j++) P = nt::mulmod(P, h->moduli[h->n_q + j] % q, q);
            }
            for (size_t c = 0; c < N; c++) {
                const uint64_t x = uniform_mod(h->rng, q);
                a[(size_t)i * N + c] = x; na[(size_t)i * N + c] = x ? q - x : 0; f[(size_t)i * N + c] = P;
            }
        }
        CK(cudaMemcpy(d_a, a.data(), sizeof(uint64_t) * a.size(), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_na, na.data(), sizeof(uint64_t) * a.size(), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_f, f.data(), sizeof(uint64_t) * a.size(), cudaMemcpyHostToDevice));
        RC(upload_signed_ntt(h, L, sample_noise(h, (int64_t)h->t), 0, total, d_e));
        RC(mac(L, {{d_na, L.d_s}, {d_e, L.d_ones}, {d_f, d_from}}, d_b, 0, total));
        uint64_t* kq = k.q + (size_t)d * 2 * l * N;
        uint64_t* kp = k.p + (size_t)d * 2 * np * N;
        CK(cudaMemcpy(kq, d_b, sizeof(uint64_t) * l * N, cudaMemcpyDeviceToDevice));
        CK(cudaMemcpy(kq + (size_t)l * N, d_a, sizeof(uint64_t) * l * N, cudaMemcpyDeviceToDevice));
        CK(cudaMemcpy(kp, d_b + (size_t)l * N, sizeof(uint64_t) * np * N, cudaMemcpyDeviceToDevice));
        CK(cudaMemcpy(kp + (size_t)np * N, d_a + (size_t)l * N, sizeof(uint64_t) * np * N, cudaMemcpyDeviceToDevice));
    }
    *out = k;
    return 0;
}

// key for source 0 (s^2) or a Galois element g (s(x^g)), made on first use
static int get_evk(b200fhe_bgv* h, Level& L, uint32_t source, EvKey* out) {
    auto it = L.keys.find(source);
    if (it != L.keys.end()) { *out = it->second; return 0; }
    if (!h->have_sk) return b200_set_err(B200FHE_EINVAL, "bgv: no secret key");
    const size_t total = (size_t)L.l + h->n_p;
    uint64_t* d_from = h->d_tmp + 5 * (size_t)(h->n_q + h->n_p) * h->N;
    if (source == 0) RC(mac(L, {{L.d_s, L.d_s}}, d_from, 0, (int)total));
    else RC(b200fhe_galois_ntt(L.ctx, L.d_s, d_from, source, total, nullptr));
    EvKey k;
    RC(make_evk(h, L, d_from, &k));
    L.keys[source] = k;
    *out = k;
    return 0;
}

// ---- batch encoder ----------------------------------------------------------------------------------------------------
// PhantomBatchEncoder::encode (dotprod_test.cu:100-101): values (< t, at most N of them, the rest zero) -> plaintext
// coefficients mod t: slot i goes to evaluation point index_map[i], then one inverse NTT modulo t (on the device).
extern "C" int b200fhe_bgv_encode(b200fhe_bgv* h, const uint64_t* values, size_t n_values, uint64_t* plain) {
    if (!h || !values || !plain || n_values > h->N) return b200_set_err(B200FHE_EINVAL, "bgv_encode: bad arguments");
    std::vector<uint64_t> slots(h->N, 0);
    for (size_t i = 0; i < n_values; i++) slots[h->index_map[i]] = values[i] % h->t;
    RC(b200fhe_nwt_backward_host(h->tctx, slots.data(), 1, 1, 0));
    for (size_t i = 0; i < h->N; i++) plain[i] = slots[i];
    return 0;
}
// PhantomBatchEncoder::decode (:119,153)
extern "C" int b200fhe_bgv_decode(b200fhe_bgv* h, const uint64_t* plain, uint64_t* values) {
    if (!h || !values || !plain) return b200_set_err(B200FHE_EINVAL, "bgv_decode: bad arguments");
    std::vector<uint64_t> slots(plain, plain + h->N);
    RC(b200fhe_nwt_forward_host(h->tctx, slots.data(), 1, 1, 0));
    for (size_t i = 0; i < h->N; i++) values[i] = slots[h->index_map[i]];
    return 0;
}

// ---- encryption / decryption ------------------------------------------------------------------------------------------
// PhantomPublicKey::encrypt_asymmetric (:105-106): d_ct [2][n_q][N] (NTT form, the Phantom device layout :42-46)
//   c0 = pk0 u + t e0 + m,  c1 = pk1 u + t e1.
extern "C" int b200fhe_bgv_encrypt(b200fhe_bgv* h, const uint64_t* plain, uint64_t* d_ct) {
    if (!h || !plain || !d_ct) return b200_set_err(B200FHE_EINVAL, "bgv_encrypt: null argument");
    if (!h->have_pk) return b200_set_err(B200FHE_EINVAL, "bgv_encrypt: no public key");
    Level* L;
    RC(ensure_level(h, h->n_q, &L));
    const size_t N = h->N, nq = (size_t)h->n_q;
    uint64_t *d_u = h->d_tmp, *d_e = d_u + nq * N, *d_m = d_e + nq * N;
    std::vector<int64_t> m(N);
    for (size_t i = 0; i < N; i++) m[i] = (int64_t)(plain[i] % h->t);
    RC(upload_signed_ntt(h, *L, sample_ternary(h), 0, (int)nq, d_u));
    RC(upload_signed_ntt(h, *L, m, 0, (int)nq, d_m));
    RC(upload_signed_ntt(h, *L, sample_noise(h, (int64_t)h->t), 0, (int)nq, d_e));
    RC(mac(*L, {{h->d_pk, d_u}, {d_e, L->d_ones}, {d_m, L->d_ones}}, d_ct, 0, (int)nq));
    RC(upload_signed_ntt(h, *L, sample_noise(h, (int64_t)h->t), 0, (int)nq, d_e));
    RC(mac(*L, {{h->d_pk + nq * N, d_u}, {d_e, L->d_ones}}, d_ct + nq * N, 0, (int)nq));
    CK(cudaDeviceSynchronize());
    return 0;
}

// PhantomSecretKey::decrypt (:118,152): d_ct [size][n_limbs][N], size 2 or 3 -> plaintext coefficients mod t.
// phase = c0 + c1 s (+ c2 s^2) on the device, inverse NTT, then per coefficient the residues are recombined exactly (Garner
// mixed radix, compared digit by digit with (Q - 1) / 2 for the centred lift) and reduced mod t; `correction` is the
// ciphertext's correction factor (the decrypted value is m * correction mod t): its inverse is multiplied in.
extern "C" int b200fhe_bgv_decrypt(b200fhe_bgv* h, const uint64_t* d_ct, int size, int n_limbs, uint64_t correction, uint64_t* plain) {
    if (!h || !d_ct || !plain || size < 2 || size > 3) return b200_set_err(B200FHE_EINVAL, "bgv_decrypt: bad arguments");
    if (!h->have_sk) return b200_set_err(B200FHE_EINVAL, "bgv_decrypt: no secret key");
    Level* L;
    RC(ensure_level(h, n_limbs, &L));
    const size_t N = h->N;
    const int l = n_limbs;
    uint64_t *d_ph = h->d_tmp, *d_s2 = d_ph + (size_t)l * N;
    if (size == 3) {
        RC(mac(*L, {{L->d_s, L->d_s}}, d_s2, 0, l));
        RC(mac(*L, {{d_ct, L->d_ones}, {d_ct + (size_t)l * N, L->d_s}, {d_ct + 2 * (size_t)l * N, d_s2}}, d_ph, 0, l));
    } else {
        RC(mac(*L, {{d_ct, L->d_ones}, {d_ct + (size_t)l * N, L->d_s}}, d_ph, 0, l));
    }
    RC(b200fhe_nwt_backward(L->ctx, d_ph, 1, (size_t)l, 0, nullptr));
    std::vector<uint64_t> r((size_t)l * N);
    CK(cudaMemcpy(r.data(), d_ph, sizeof(uint64_t) * r.size(), cudaMemcpyDeviceToHost));
    const uint64_t t = h->t;
    // Garner constants: inv[i][j] = q_j^-1 mod q_i (j < i); radix weights mod t; Q mod t
    std::vector<std::vector<uint64_t>> inv(l);
    std::vector<uint64_t> w(l), half(l);
    uint64_t acc = 1 % t;
    for (int i = 0; i < l; i++) {
        const uint64_t qi = h->moduli[i];
        inv[i].resize(i);
        for (int j = 0; j < i; j++) inv[i][j] = nt::invmod(h->moduli[j] % qi, qi);
        w[i] = acc;
        acc = nt::mulmod(acc, qi % t, t);
        half[i] = (qi - 1) / 2;
    }
    const uint64_t Q_mod_t = acc;
    const uint64_t cinv = nt::invmod(correction % t ? correction % t : 1, t);
    std::vector<uint64_t> d(l);
    for (size_t c = 0; c < N; c++) {
        for (int i = 0; i < l; i++) {
            const uint64_t qi = h->moduli[i];
            uint64_t v = r[(size_t)i * N + c] % qi;
            for (int j = 0; j < i; j++) {                       // v = (v - d_j) q_j^-1 mod q_i
                const uint64_t dj = d[j] % qi;
                v = nt::mulmod(v >= dj ? v - dj : v + qi - dj, inv[i][j], qi);
            }
            d[i] = v;
        }
        bool neg = false;                                       // x > (Q - 1) / 2 ?
        for (int i = l - 1; i >= 0; i--)
            if (d[i] != half[i]) { neg = d[i] > half[i]; break; }
        uint64_t x = 0;
        for (int i = 0; i < l; i++) x = (x + nt::mulmod(d[i] % t, w[i], t)) % t;
        if (neg) x = (x + t - Q_mod_t) % t;
        plain[c] = nt::mulmod(x, cinv, t);
    }
    return 0;
}

// ---- evaluator --------------------------------------------------------------------------------------------------------
// phantom::multiply (:113): (a0, a1) x (b0, b1) -> (a0 b0, a0 b1 + a1 b0, a1 b1); d_out [3][n_limbs][N]
extern "C" int b200fhe_bgv_multiply(b200fhe_bgv* h, const uint64_t* d_a, const uint64_t* d_b, uint64_t* d_out, int n_limbs) {
    if (!h) return b200_set_err(B200FHE_EINVAL, "bgv_multiply: null");
    Level* L;
    RC(ensure_level(h, n_limbs, &L));
    return b200fhe_ct_tensor(L->ctx, d_a, d_b, d_out, (size_t)n_limbs, 0, nullptr);
}

// phantom::relinearize_inplace (:114): d_ct [3][n_limbs][N] -> its first two polynomials become (d0, d1) + KeySwitch(d2)
extern "C" int b200fhe_bgv_relinearize(b200fhe_bgv* h, uint64_t* d_ct, int n_limbs) {
    if (!h || !d_ct) return b200_set_err(B200FHE_EINVAL, "bgv_relinearize: null");
    Level* L;
    RC(ensure_level(h, n_limbs, &L));
    EvKey k;
    RC(get_evk(h, *L, 0, &k));
    const size_t N = h->N, l = (size_t)n_limbs;
    uint64_t* d_sw = h->d_tmp + 6 * (size_t)(h->n_q + h->n_p) * N;
    RC(b200fhe_keyswitch_apply(L->ks, d_ct + 2 * l * N, k.q, k.p, d_sw, nullptr));
    return b200fhe_poly_add(L->ctx, d_ct, d_sw, 2, l, 0, nullptr);
}

// phantom::mod_switch_to_next_inplace (:115): d_ct [size][n_limbs][N] -> d_out [size][n_limbs - 1][N] = (ct - t [ct t^-1]_q_last) / q_last;
// the plaintext becomes m q_last^-1 mod t: *correction (in/out) is multiplied by q_last^-1 mod t.
extern "C" int b200fhe_bgv_mod_switch_to_next(b200fhe_bgv* h, const uint64_t* d_ct, int size, int n_limbs, uint64_t* d_out, uint64_t* correction) {
    if (!h || !d_ct || !d_out || size < 1 || size > 3) return b200_set_err(B200FHE_EINVAL, "bgv_mod_switch_to_next: bad arguments");
    if (n_limbs < 2) return b200_set_err(B200FHE_EINVAL, "bgv_mod_switch_to_next: already at the last level");
    Level* L;
    RC(ensure_level(h, n_limbs, &L));
    const size_t N = h->N, l = (size_t)n_limbs;
    uint64_t *d_q = h->d_tmp, *d_p = d_q + 2 * l * N;
    for (int p0 = 0; p0 < size; p0 += 2) {
        const int n = size - p0 < 2 ? size - p0 : 2;
        for (int p = 0; p < n; p++) {
            CK(cudaMemcpyAsync(d_q + (size_t)p * (l - 1) * N, d_ct + (size_t)(p0 + p) * l * N, sizeof(uint64_t) * (l - 1) * N, cudaMemcpyDeviceToDevice, nullptr));
            CK(cudaMemcpyAsync(d_p + (size_t)p * N, d_ct + ((size_t)(p0 + p) * l + (l - 1)) * N, sizeof(uint64_t) * N, cudaMemcpyDeviceToDevice, nullptr));
        }
        RC(b200fhe_keyswitch_moddown(L->down, d_q, d_p, d_out + (size_t)p0 * (l - 1) * N, (size_t)n, nullptr));
    }
    if (correction) *correction = nt::mulmod(*correction % h->t, nt::invmod(h->moduli[n_limbs - 1] % h->t, h->t), h->t);
    return 0;
}

// phantom::rotate_inplace (:146): rows rotated left by `step` (Galois element 3^step mod 2N; step < 0: the inverse element;
// step = 0 with columns = 1: the column swap, element 2N - 1); d_ct [2][n_limbs][N] in place.
extern "C" int b200fhe_bgv_rotate(b200fhe_bgv* h, uint64_t* d_ct, int n_limbs, int step, int columns) {
    if (!h || !d_ct) return b200_set_err(B200FHE_EINVAL, "bgv_rotate: null");
    Level* L;
    RC(ensure_level(h, n_limbs, &L));
    const uint64_t two_n = 2 * h->N, row = h->N / 2;
    uint32_t g;
    if (columns) g = (uint32_t)(two_n - 1);
    else {
        const uint64_t e = (uint64_t)(((int64_t)step % (int64_t)row + (int64_t)row) % (int64_t)row);
        if (!e) return 0;
        g = (uint32_t)nt::powmod(kGen, e, two_n);
    }
    EvKey k;
    RC(get_evk(h, *L, g, &k));
    const size_t N = h->N, l = (size_t)n_limbs;
    uint64_t *d_g = h->d_tmp + 6 * (size_t)(h->n_q + h->n_p) * N;       // [2][l][N] permuted ciphertext
    RC(b200fhe_galois_ntt(L->ctx, d_ct, d_g, g, 2 * l, nullptr));
    RC(b200fhe_keyswitch_apply(L->ks, d_g + l * N, k.q, k.p, d_ct, nullptr));
    return b200fhe_poly_add(L->ctx, d_ct, d_g, 1, l, 0, nullptr);
}

// phantom::add_inplace (:150)
extern "C" int b200fhe_bgv_add(b200fhe_bgv* h, uint64_t* d_a, const uint64_t* d_b, int size, int n_limbs) {
    if (!h) return b200_set_err(B200FHE_EINVAL, "bgv_add: null");
    Level* L;
    RC(ensure_level(h, n_limbs, &L));
    return b200fhe_poly_add(L->ctx, d_a, d_b, (size_t)size, (size_t)n_limbs, 0, nullptr);
}
