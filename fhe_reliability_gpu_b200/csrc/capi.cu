// C ABI (include/b200fhe.h): context management and the NTT entry points.
#include "../../include/b200fhe.h"
#include "ctx.hpp"
#include "numth.hpp"
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <atomic>

using namespace b200;

static thread_local char g_err[512] = "";

int b200_set_err(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return b200_set_err(B200FHE_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                                __FILE__, __LINE__);                                               \
    } while (0)

extern "C" int b200fhe_version(void) { return 100; }
extern "C" const char* b200fhe_last_error(void) { return g_err; }

extern "C" int b200fhe_primes_create(uint64_t N, const int* bits, int n, uint64_t* out) {
    if (!bits || !out || n < 0 || N < 2 || (N & (N - 1))) return b200_set_err(B200FHE_EINVAL, "primes_create: bad args");
    if (!nt::primes_create(N, bits, n, out)) return b200_set_err(B200FHE_EINVAL, "primes_create: not enough primes");
    return 0;
}
extern "C" int b200fhe_min_primitive_root(uint64_t N, uint64_t q, uint64_t* psi) {
    if (!psi) return b200_set_err(B200FHE_EINVAL, "null psi");
    uint64_t r = nt::min_primitive_root(2 * N, q);
    if (!r) return b200_set_err(B200FHE_EINVAL, "no primitive 2N-th root: q=%llu N=%llu", (unsigned long long)q, (unsigned long long)N);
    *psi = r;
    return 0;
}
extern "C" int b200fhe_host_tables(int log_n, uint64_t q, uint64_t* rp, uint64_t* rps, uint64_t* irp, uint64_t* irps,
                                   uint64_t* n_inv, uint64_t* n_inv_shoup) {
    if (log_n < 1 || log_n > 20) return b200_set_err(B200FHE_EINVAL, "log_n out of range");
    uint64_t psi = nt::min_primitive_root((uint64_t)2 << log_n, q);
    if (!psi) return b200_set_err(B200FHE_EINVAL, "modulus %llu has no primitive 2N-th root", (unsigned long long)q);
    nt::tables(log_n, q, psi, rp, rps, irp, irps);
    uint64_t ni = nt::invmod(((uint64_t)1 << log_n) % q, q);
    if (n_inv) *n_inv = ni;
    if (n_inv_shoup) *n_inv_shoup = nt::shoup(ni, q);
    return 0;
}
extern "C" int b200fhe_host_tables_psi(int log_n, uint64_t q, uint64_t psi, uint64_t* rp, uint64_t* rps, uint64_t* irp,
                                       uint64_t* irps, uint64_t* n_inv, uint64_t* n_inv_shoup) {
    if (log_n < 1 || log_n > 20) return b200_set_err(B200FHE_EINVAL, "log_n out of range");
    if (q < 3 || (q >> 60) || psi == 0 || psi >= q) return b200_set_err(B200FHE_EINVAL, "bad modulus or root");
    uint64_t t = psi;
    for (int i = 0; i < log_n; i++) t = nt::mulmod(t, t, q);          // psi^N must be -1
    if (t != q - 1) return b200_set_err(B200FHE_EINVAL, "psi=%llu is not a primitive 2N-th root mod %llu", (unsigned long long)psi, (unsigned long long)q);
    nt::tables(log_n, q, psi, rp, rps, irp, irps);
    uint64_t ni = nt::invmod(((uint64_t)1 << log_n) % q, q);
    if (n_inv) *n_inv = ni;
    if (n_inv_shoup) *n_inv_shoup = nt::shoup(ni, q);
    return 0;
}
extern "C" int b200fhe_const_ratio(uint64_t q, uint64_t ratio[2]) {
    if (q < 2 || !ratio) return b200_set_err(B200FHE_EINVAL, "bad modulus");
    nt::const_ratio(q, ratio);
    return 0;
}

// ---------------------------------------------------------------- context
B200Mod b200_make_mod(uint64_t q, uint64_t itw1, uint64_t n_inv) {
    B200Mod m;
    memset(&m, 0, sizeof(m));
    m.q = q;
    m.bmu = ~(uint64_t)0 / q;
    m.ninv = n_inv;
    m.ninv_s = nt::shoup(n_inv, q);
    m.iw1n = nt::mulmod(itw1, n_inv, q);
    m.iw1n_s = nt::shoup(m.iw1n, q);
    uint64_t r[2];
    nt::const_ratio(q, r);
    m.ratio0 = r[0];
    m.ratio1 = r[1];
    m.bits = nt::bitlen(q);
    m.big = m.bits > 36;
    m.mprec = m.big ? (uint32_t)((((nt::u128)1) << (m.bits + 31)) / q) : 0;
    m.pshift = m.big ? m.bits - 33 : 0;
    m.q3 = 3 * q;
    const uint64_t delta = m.bits < 64 ? ((uint64_t)1 << m.bits) - q : 0;
    m.near = m.big && m.bits >= 40 && delta < ((uint64_t)1 << 31) &&
             ((nt::u128)delta << (64 - m.bits)) < (q >> 1);
    m.delta = m.near ? (uint32_t)delta : 0;
    return m;
}

extern "C" int b200fhe_ctx_create_empty(int log_n, int L, int device, b200fhe_ctx** out) {
    if (!out) return b200_set_err(B200FHE_EINVAL, "null out");
    if (log_n < 1 || log_n > 17) return b200_set_err(B200FHE_EUNSUPPORTED, "log_n=%d unsupported (1..17)", log_n);
    if (L < 1 || L > 4096) return b200_set_err(B200FHE_EINVAL, "L=%d out of range", L);
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return b200_set_err(B200FHE_ECUDA, "no CUDA device: this library has no CPU fallback");
    if (device < 0 || device >= ndev) return b200_set_err(B200FHE_EINVAL, "device %d of %d", device, ndev);
    CK(cudaSetDevice(device));
    b200fhe_ctx* c = new b200fhe_ctx();
    c->log_n = log_n;
    c->N = (size_t)1 << log_n;
    c->L = L;
    c->device = device;
    c->moduli.assign(L, 0);
    c->psi.assign(L, 0);
    c->h_mods.resize(L);
    memset(c->h_mods.data(), 0, sizeof(B200Mod) * L);
    c->has_inverse.assign(L, 0);
    c->has_omega.assign(L, 0);
    cudaError_t e = cudaMalloc(&c->d_mods, sizeof(B200Mod) * L);
    if (e == cudaSuccess) e = cudaMalloc(&c->d_tw, sizeof(ulonglong2) * c->N * L);
    if (e == cudaSuccess) e = cudaMalloc(&c->d_itw, sizeof(ulonglong2) * c->N * L);
    if (e == cudaSuccess && log_n >= 12) e = cudaMalloc(&c->d_rtw, sizeof(ulonglong2) * c->N * L);
    if (e == cudaSuccess && log_n >= 12) e = cudaMalloc(&c->d_irtw, sizeof(ulonglong2) * c->N * L);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        b200fhe_ctx_destroy(c);
        return b200_set_err(e == cudaErrorMemoryAllocation ? B200FHE_ENOMEM : B200FHE_ECUDA, "ctx alloc: %s", cudaGetErrorString(e));
    }
    *out = c;
    return 0;
}

// Row-pass tiles (ntt_kernels.cu RowSmem): entries [N/256, N) of a SEAL-layout
// table regrouped per tile of R = kRowsPerCta rows as A[row][15+pad] (stages
// 0..3 of the row) followed by B[j][thread] (stages 4..7): 256 R entries per tile.
static void build_row_tiles(const std::vector<ulonglong2>& tw, size_t N, std::vector<ulonglong2>& out) {
    const size_t CL = N >> 8, R = b200::kRowsPerCta, TH = R * 16;
    out.assign(N, make_ulonglong2(0, 0));
    for (size_t T = 0; T < CL / R; T++) {
        ulonglong2* tile = out.data() + T * b200::kRowTileEntries;
        for (size_t rl = 0; rl < R; rl++) {
            const size_t P = CL + R * T + rl;
            for (int t = 0; t < 4; t++)
                for (size_t g = 0; g < ((size_t)1 << t); g++) {
                    const size_t j = ((size_t)1 << t) - 1 + g;
                    tile[rl * 16 + j] = tw[(P << t) + g];
                    for (size_t cl = 0; cl < 16; cl++)
                        tile[R * 16 + j * TH + rl * 16 + cl] = tw[((((P << 4) + cl)) << t) + g];
                }
        }
    }
}

extern "C" int b200fhe_ctx_set_tables(b200fhe_ctx* c, int idx, uint64_t q, const uint64_t* tw, const uint64_t* tws,
                                      const uint64_t* itw, const uint64_t* itws, uint64_t n_inv, uint64_t n_inv_shoup,
                                      void* stream) {
    if (!c || idx < 0 || idx >= c->L || !tw || !tws) return b200_set_err(B200FHE_EINVAL, "set_tables: bad args");
    if (q < 3 || (q >> 60)) return b200_set_err(B200FHE_EINVAL, "modulus %llu outside (2, 2^60)", (unsigned long long)q);
    (void)n_inv_shoup;
    CK(cudaSetDevice(c->device));
    cudaStream_t s = (cudaStream_t)stream;
    const size_t N = c->N;
    std::vector<ulonglong2> buf(N);
    for (size_t i = 0; i < N; i++) buf[i] = make_ulonglong2(tw[i], tws[i]);
    CK(cudaMemcpyAsync(c->d_tw + (size_t)idx * N, buf.data(), sizeof(ulonglong2) * N, cudaMemcpyHostToDevice, s));
    CK(cudaStreamSynchronize(s));
    std::vector<ulonglong2> tiles;
    if (c->d_rtw) {
        build_row_tiles(buf, N, tiles);
        CK(cudaMemcpyAsync(c->d_rtw + (size_t)idx * N, tiles.data(), sizeof(ulonglong2) * N, cudaMemcpyHostToDevice, s));
        CK(cudaStreamSynchronize(s));
    }
    uint64_t itw1 = 0;
    if (itw && itws) {
        for (size_t i = 0; i < N; i++) buf[i] = make_ulonglong2(itw[i], itws[i]);
        CK(cudaMemcpyAsync(c->d_itw + (size_t)idx * N, buf.data(), sizeof(ulonglong2) * N, cudaMemcpyHostToDevice, s));
        CK(cudaStreamSynchronize(s));
        if (c->d_irtw) {
            build_row_tiles(buf, N, tiles);
            CK(cudaMemcpyAsync(c->d_irtw + (size_t)idx * N, tiles.data(), sizeof(ulonglong2) * N, cudaMemcpyHostToDevice, s));
            CK(cudaStreamSynchronize(s));
        }
        itw1 = N > 1 ? itw[1] : 1;
        c->has_inverse[idx] = 1;
        if (!n_inv) n_inv = nt::invmod(N % q, q);
    } else {
        c->has_inverse[idx] = 0;
    }
    c->moduli[idx] = q;
    c->psi[idx] = N > 1 ? tw[N >> 1] : 1;   // root_powers[bitrev(1)] = psi
    c->has_omega[idx] = 0;
    c->h_mods[idx] = b200_make_mod(q, itw1, n_inv);
    CK(cudaMemcpyAsync(c->d_mods + idx, &c->h_mods[idx], sizeof(B200Mod), cudaMemcpyHostToDevice, s));
    CK(cudaStreamSynchronize(s));
    return 0;
}

extern "C" int b200fhe_ctx_create(int log_n, const uint64_t* moduli, int L, int device, b200fhe_ctx** out) {
    if (!moduli) return b200_set_err(B200FHE_EINVAL, "null moduli");
    const uint64_t N = (uint64_t)1 << log_n;
    for (int i = 0; i < L; i++) {
        uint64_t q = moduli[i];
        if (q < 3 || (q >> 60) || (q - 1) % (2 * N) || !nt::is_prime(q))
            return b200_set_err(B200FHE_EINVAL, "modulus[%d]=%llu must be a prime < 2^60 with q = 1 mod 2N", i, (unsigned long long)q);
    }
    int rc = b200fhe_ctx_create_empty(log_n, L, device, out);
    if (rc) return rc;
    std::vector<uint64_t> rp(N), rps(N), irp(N), irps(N);
    for (int i = 0; i < L; i++) {
        uint64_t q = moduli[i], psi = nt::min_primitive_root(2 * N, q);
        if (!psi) { b200fhe_ctx_destroy(*out); *out = nullptr; return b200_set_err(B200FHE_EINVAL, "no root for modulus %d", i); }
        nt::tables(log_n, q, psi, rp.data(), rps.data(), irp.data(), irps.data());
        uint64_t ni = nt::invmod(N % q, q);
        rc = b200fhe_ctx_set_tables(*out, i, q, rp.data(), rps.data(), irp.data(), irps.data(), ni, nt::shoup(ni, q), nullptr);
        if (rc) { b200fhe_ctx_destroy(*out); *out = nullptr; return rc; }
        (*out)->psi[i] = psi;
    }
    return 0;
}

extern "C" int b200fhe_ctx_destroy(b200fhe_ctx* c) {
    if (!c) return 0;
    cudaSetDevice(c->device);
    if (c->d_mods) cudaFree(c->d_mods);
    if (c->d_tw) cudaFree(c->d_tw);
    if (c->d_itw) cudaFree(c->d_itw);
    if (c->d_rtw) cudaFree(c->d_rtw);
    if (c->d_irtw) cudaFree(c->d_irtw);
    if (c->d_omega) cudaFree(c->d_omega);
    if (c->d_abft) cudaFree(c->d_abft);
    if (c->d_inj) cudaFree(c->d_inj);
    if (c->d_stage) cudaFree(c->d_stage);
    if (c->h_stage) cudaFreeHost(c->h_stage);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    if (c->s_h2d) cudaStreamDestroy(c->s_h2d);
    if (c->s_d2h) cudaStreamDestroy(c->s_d2h);
    for (auto e : c->ev) cudaEventDestroy(e);
    delete c;
    return 0;
}
extern "C" int b200fhe_ctx_info(const b200fhe_ctx* c, int* log_n, int* L, int* device) {
    if (!c) return b200_set_err(B200FHE_EINVAL, "null ctx");
    if (log_n) *log_n = c->log_n;
    if (L) *L = c->L;
    if (device) *device = c->device;
    return 0;
}
extern "C" int b200fhe_ctx_moduli(const b200fhe_ctx* c, uint64_t* out) {
    if (!c || !out) return b200_set_err(B200FHE_EINVAL, "null arg");
    memcpy(out, c->moduli.data(), sizeof(uint64_t) * c->L);
    return 0;
}

// ---------------------------------------------------------------- transforms
static std::atomic<unsigned long long> g_launches{0};   // kernels launched by this library (bench.py's gpu_launches)
extern "C" unsigned long long b200fhe_launch_count(void) { return g_launches.load(); }
void b200_count_launches(unsigned n) { g_launches += n; }

// Polynomials per launch on the device-resident path.  The transform is bound by the integer
// multiplier pipe, not by HBM (profiles/r1e_*), so keeping the lazily-reduced intermediate inside
// the 126 MB L2 buys nothing while short launches lose 13 % to wave tails (chunk sweep in
// profiles/r1e_chunk_sweep.md): the whole batch goes out as one launch pair per 32768 limbs.
static size_t chunk_polys(const b200fhe_ctx* c, size_t L) {
    static const size_t budget = [] {
        const char* e = getenv("B200FHE_CHUNK_MB");
        return (size_t)(e && atoi(e) > 0 ? atoi(e) : 1 << 20) << 20;
    }();
    const size_t bytes_per_poly = L * c->N * sizeof(uint64_t);
    size_t n = budget / (bytes_per_poly ? bytes_per_poly : 1);
    const size_t cap = 32768 / (L ? L : 1);          // grid.y limit of one launch
    if (n > cap) n = cap;
    return n ? n : 1;
}
static int run_ntt(b200fhe_ctx* c, bool inverse, uint64_t* d, size_t n_poly, size_t L, size_t start, int mask,
                   cudaStream_t s, const B200Mod* mods_alt = nullptr, const uint64_t* src = nullptr) {
    if (!c || (!d && n_poly && L)) return b200_set_err(B200FHE_EINVAL, "nwt: null argument");
    if (L == 0 || n_poly == 0) return 0;
    if (start + L > (size_t)c->L) return b200_set_err(B200FHE_EINVAL, "nwt: start_mod %zu + L %zu exceeds context (%d)", start, L, c->L);
    bool all_big = true, all_near = true;
    for (size_t i = start; i < start + L; i++) {
        if (!c->moduli[i]) return b200_set_err(B200FHE_EINVAL, "nwt: modulus %zu not set", i);
        if (inverse && !c->has_inverse[i]) return b200_set_err(B200FHE_EINVAL, "nwt: no inverse tables for modulus %zu", i);
        all_big = all_big && c->h_mods[i].big;
        all_near = all_near && c->h_mods[i].near;
    }
    static const bool no_near = getenv("B200FHE_NO_NEAR") != nullptr;     // A/B switch for the NEAR kernels
    const int big = (all_near && !no_near) ? 2 : all_big ? 1 : 0;
    CK(cudaSetDevice(c->device));
    PassArgs a;
    a.mods = mods_alt ? mods_alt : c->d_mods;
    a.tw = inverse ? c->d_itw : c->d_tw;
    a.rtw = inverse ? c->d_irtw : c->d_rtw;
    a.N = (u32)c->N;
    a.L = (u32)L;
    a.start_mod = (u32)start;
    const size_t step = chunk_polys(c, L);
    for (size_t p0 = 0; p0 < n_poly; p0 += step) {
        const size_t np = n_poly - p0 < step ? n_poly - p0 : step;
        a.data = (u64*)d + p0 * L * c->N;
        a.src = src ? (const u64*)src + p0 * L * c->N : nullptr;
        int rc = launch_ntt(inverse, big, a, c->log_n, (u32)(np * L), mask, s);
        if (rc) return b200_set_err(B200FHE_EUNSUPPORTED, "nwt: log_n=%d unsupported (rc=%d)", c->log_n, rc);
        g_launches += (c->log_n <= 11) ? 1 : ((mask == 3) ? 2 : 1);
    }
    CK(cudaGetLastError());
    return 0;
}

extern "C" int b200fhe_nwt_forward(b200fhe_ctx* c, uint64_t* d, size_t n_poly, size_t L, size_t start, void* stream) {
    return run_ntt(c, false, d, n_poly, L, start, 3, (cudaStream_t)stream);
}
extern "C" int b200fhe_nwt_backward(b200fhe_ctx* c, uint64_t* d, size_t n_poly, size_t L, size_t start, void* stream) {
    return run_ntt(c, true, d, n_poly, L, start, 3, (cudaStream_t)stream);
}
// Inverse transform whose final N^-1 scaling uses an alternative per-modulus constant table (same layout as the context's
// B200Mod array, ninv / iw1n replaced by ninv c_l / iw1n c_l): the key-switch folds the base conversion's scale step
// y_j = x_j (D/q_j)^-1 into the inverse NTT this way (keyswitch.cu).
int b200_nwt_backward_mods(b200fhe_ctx* c, uint64_t* d, size_t n_poly, size_t L, size_t start, const B200Mod* d_mods_alt, void* stream) {
    return run_ntt(c, true, d, n_poly, L, start, 3, (cudaStream_t)stream, d_mods_alt);
}
// the same, out of place: d = INTT(d_src) (scaled), d_src untouched.  Sizes with the two-pass kernels only (log_n >= 12): returns 1
// otherwise and the caller copies first.
int b200_nwt_backward_mods_oop(b200fhe_ctx* c, const uint64_t* d_src, uint64_t* d, size_t n_poly, size_t L, size_t start, const B200Mod* d_mods_alt,
                               void* stream) {
    if (!c || c->log_n <= 11 || !d_src) return 1;
    return run_ntt(c, true, d, n_poly, L, start, 3, (cudaStream_t)stream, d_mods_alt, d_src);
}
// Forward transform of n scattered limbs in ONE launch pair: entry y = data limb h_map[y].x (offset in limbs from d) with modulus
// h_map[y].y; d_map is the same list on the device.  Returns 1 (not an error) when this size has no scatter form (log_n <= 11):
// the caller falls back to range launches.  Used by the key-switch for the converted limbs of all digits (keyswitch.cu).
int b200_nwt_forward_map(b200fhe_ctx* c, uint64_t* d, const uint2* d_map, const uint2* h_map, size_t n, void* stream) {
    if (!c || !d || !d_map || !h_map) return b200_set_err(B200FHE_EINVAL, "nwt_map: null argument");
    if (n == 0) return 0;
    if (c->log_n <= 11 || n > 65535) return 1;
    bool all_big = true, all_near = true;
    for (size_t i = 0; i < n; i++) {
        const size_t mi = h_map[i].y;
        if (mi >= (size_t)c->L || !c->moduli[mi]) return b200_set_err(B200FHE_EINVAL, "nwt_map: modulus %zu not set", mi);
        all_big = all_big && c->h_mods[mi].big;
        all_near = all_near && c->h_mods[mi].near;
    }
    static const bool no_near = getenv("B200FHE_NO_NEAR") != nullptr;
    const int big = (all_near && !no_near) ? 2 : all_big ? 1 : 0;
    CK(cudaSetDevice(c->device));
    PassArgs a;
    a.mods = c->d_mods;
    a.tw = c->d_tw;
    a.rtw = c->d_rtw;
    a.N = (u32)c->N;
    a.L = (u32)n;
    a.start_mod = 0;
    a.map = d_map;
    a.data = (u64*)d;
    const int rc = launch_ntt(false, big, a, c->log_n, (u32)n, 3, (cudaStream_t)stream);
    if (rc) return b200_set_err(B200FHE_EUNSUPPORTED, "nwt_map: log_n=%d unsupported (rc=%d)", c->log_n, rc);
    g_launches += 2;
    CK(cudaGetLastError());
    return 0;
}
// measurement hook: one pass (1 = first kernel, 2 = second kernel) of the transform
extern "C" int b200fhe_debug_nwt_pass(b200fhe_ctx* c, uint64_t* d, size_t n_poly, size_t L, size_t start, int inverse,
                                      int pass, void* stream) {
    if (pass != 1 && pass != 2) return b200_set_err(B200FHE_EINVAL, "pass must be 1 or 2");
    return run_ntt(c, inverse != 0, d, n_poly, L, start, pass, (cudaStream_t)stream);
}

int b200_ensure_stage(b200fhe_ctx* c, size_t words, bool need_host) {
    if (c->stage_words < words) {
        if (c->d_stage) cudaFree(c->d_stage);
        c->d_stage = nullptr; c->stage_words = 0;
        CK(cudaMalloc(&c->d_stage, words * sizeof(uint64_t)));
        c->stage_words = words;
    }
    if (need_host && c->hstage_words < words) {
        if (c->h_stage) cudaFreeHost(c->h_stage);
        c->h_stage = nullptr; c->hstage_words = 0;
        CK(cudaMallocHost(&c->h_stage, words * sizeof(uint64_t)));
        c->hstage_words = words;
    }
    return 0;
}

extern "C" int b200fhe_host_alloc(void** p, size_t bytes) {
    if (!p) return b200_set_err(B200FHE_EINVAL, "null out");
    CK(cudaMallocHost(p, bytes));
    return 0;
}
extern "C" int b200fhe_host_free(void* p) {
    if (p) CK(cudaFreeHost(p));
    return 0;
}

static bool is_pinned(const void* p) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeHost;
}

// H2D -> kernels -> D2H as in ntt_test.cu:89-102, but chunked per polynomial
// group and software-pipelined over three streams so the two PCIe directions
// and the kernels overlap.  Pinned caller buffers are copied directly; pageable
// ones go through the context's pinned staging buffer.
static int run_ntt_host(b200fhe_ctx* c, bool inverse, uint64_t* h, size_t n_poly, size_t L, size_t start) {
    if (!c || (!h && n_poly && L)) return b200_set_err(B200FHE_EINVAL, "nwt_host: null argument");
    if (n_poly == 0 || L == 0) return 0;
    CK(cudaSetDevice(c->device));
    const size_t poly_words = L * c->N, words = n_poly * poly_words;
    const bool pinned = is_pinned(h);
    int rc = b200_ensure_stage(c, words, !pinned);
    if (rc) return rc;
    uint64_t* src = h;
    if (!pinned) { memcpy(c->h_stage, h, words * sizeof(uint64_t)); src = c->h_stage; }
    if (!c->s_h2d) {
        CK(cudaStreamCreateWithFlags(&c->s_h2d, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&c->s_d2h, cudaStreamNonBlocking));
    }
    // pipeline pieces: whole polynomials when they are small, limb ranges of one polynomial otherwise.  The first
    // H2D and the last D2H of a call cannot overlap anything, so pieces are kept near 16 MiB
    // (B200FHE_HOST_CHUNK_MB; sweep 2/4/8/16/36 MiB -> 58/65/74/79/74 k limb-transforms/s end to end).
    struct Piece { size_t off, np, L, start; };
    std::vector<Piece> pieces;
    static const size_t target = [] {
        const char* e = getenv("B200FHE_HOST_CHUNK_MB");
        return (size_t)(e && atoi(e) > 0 ? atoi(e) : 16) << 20;
    }();
    const size_t limb_bytes = c->N * sizeof(uint64_t), poly_bytes = L * limb_bytes;
    if (poly_bytes <= target) {
        const size_t step = target / poly_bytes;
        for (size_t p0 = 0; p0 < n_poly; p0 += step)
            pieces.push_back({p0 * poly_words, n_poly - p0 < step ? n_poly - p0 : step, L, start});
    } else {
        const size_t lc = target / limb_bytes ? target / limb_bytes : 1;
        for (size_t p0 = 0; p0 < n_poly; p0++)
            for (size_t l0 = 0; l0 < L; l0 += lc)
                pieces.push_back({p0 * poly_words + l0 * c->N, 1, L - l0 < lc ? L - l0 : lc, start + l0});
    }
    const size_t n_chunks = pieces.size();
    if (c->ev.size() < 2 * n_chunks) {
        size_t old = c->ev.size();
        c->ev.resize(2 * n_chunks);
        for (size_t i = old; i < c->ev.size(); i++) CK(cudaEventCreateWithFlags(&c->ev[i], cudaEventDisableTiming));
    }
    cudaStream_t sk = c->own_stream;
    for (size_t k = 0; k < n_chunks; k++) {
        const Piece& pc = pieces[k];
        const size_t off = pc.off, cnt = pc.np * pc.L * c->N;
        CK(cudaMemcpyAsync(c->d_stage + off, src + off, cnt * sizeof(uint64_t), cudaMemcpyHostToDevice, c->s_h2d));
        CK(cudaEventRecord(c->ev[2 * k], c->s_h2d));
        CK(cudaStreamWaitEvent(sk, c->ev[2 * k], 0));
        rc = run_ntt(c, inverse, c->d_stage + off, pc.np, pc.L, pc.start, 3, sk);
        if (rc) return rc;
        CK(cudaEventRecord(c->ev[2 * k + 1], sk));
        CK(cudaStreamWaitEvent(c->s_d2h, c->ev[2 * k + 1], 0));
        CK(cudaMemcpyAsync(src + off, c->d_stage + off, cnt * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->s_d2h));
    }
    CK(cudaStreamSynchronize(c->s_d2h));
    if (!pinned) memcpy(h, c->h_stage, words * sizeof(uint64_t));
    return 0;
}
extern "C" int b200fhe_nwt_forward_host(b200fhe_ctx* c, uint64_t* h, size_t n_poly, size_t L, size_t start) {
    return run_ntt_host(c, false, h, n_poly, L, start);
}
extern "C" int b200fhe_nwt_backward_host(b200fhe_ctx* c, uint64_t* h, size_t n_poly, size_t L, size_t start) {
    return run_ntt_host(c, true, h, n_poly, L, start);
}

// ---------------------------------------------------------------- protected forward NTT
static int check_limbs(const b200fhe_ctx* c, size_t L, size_t start, const char* what);
// Counters and injection staging are rings of kAbftSlots slices, one per protected call: calls on different streams do
// not share a counter set (up to kAbftSlots in flight).  Everything is ordered on the CALLER's stream s: the counters of a
// new (or regrown) allocation are cleared with cudaMemsetAsync(s), the output weights are uploaded with
// cudaMemcpyAsync(s) + a synchronise (pageable source), so the first abft_commit RED cannot race either of them.
static int ensure_abft(b200fhe_ctx* c, size_t start, size_t L, size_t n_limbs, cudaStream_t s) {
    if (!c->d_omega) CK(cudaMalloc(&c->d_omega, sizeof(uint64_t) * c->N * c->L));
    if (!c->d_inj) CK(cudaMalloc(&c->d_inj, sizeof(InjEntry) * B200FHE_MAX_INJECT * kAbftSlots));
    if (c->abft_limbs < n_limbs) {
        if (c->d_abft) { CK(cudaDeviceSynchronize()); cudaFree(c->d_abft); }     // earlier calls may still use the old ring
        c->d_abft = nullptr; c->abft_limbs = 0;
        CK(cudaMalloc(&c->d_abft, sizeof(uint64_t) * 16 * n_limbs * kAbftSlots));
        CK(cudaMemsetAsync(c->d_abft, 0, sizeof(uint64_t) * 16 * n_limbs * kAbftSlots, s));
        CK(cudaStreamSynchronize(s));           // other streams may take the next slots
        c->abft_limbs = n_limbs;
        c->abft_dirty = 0;
    }
    std::vector<uint64_t> om;
    bool uploaded = false;
    for (size_t i = start; i < start + L; i++) {
        if (c->has_omega[i]) continue;
        om.resize(c->N);
        nt::abft_output_weights(c->log_n, c->moduli[i], c->psi[i], om.data());
        for (size_t k = 0; k < c->N; k++) om[k] = (om[k] & 0x3fffffffull) | ((om[k] >> 30) << 32);   // 30-bit halves
        CK(cudaMemcpyAsync(c->d_omega + i * c->N, om.data(), sizeof(uint64_t) * c->N, cudaMemcpyHostToDevice, s));
        CK(cudaStreamSynchronize(s));           // om is reused for the next modulus
        c->has_omega[i] = 1;
        uploaded = true;
    }
    if (uploaded) CK(cudaDeviceSynchronize());  // the table is shared by every stream
    return 0;
}

extern "C" int b200fhe_abft_weights(b200fhe_ctx* c, int idx, uint64_t* w, uint64_t* w_hat) {
    if (!c || idx < 0 || idx >= c->L || !c->moduli[idx]) return b200_set_err(B200FHE_EINVAL, "abft_weights: bad modulus index");
    if (w) for (size_t i = 0; i < c->N; i++) w[i] = nt::abft_input_weight(i, c->log_n);
    if (w_hat) nt::abft_output_weights(c->log_n, c->moduli[idx], c->psi[idx], w_hat);
    return 0;
}

__global__ void k_xor_word(u64* p, u64 mask) { *p ^= mask; }

static int nwt_protected_impl(bool inverse, b200fhe_ctx* c, uint64_t* d, size_t n_poly, size_t L, size_t start, uint32_t* d_verdict,
                              uint64_t* d_sums, const b200fhe_inject* plan, size_t n_plan, void* stream) {
    int rc = check_limbs(c, L, start, "nwt_protected");
    if (rc) return rc;
    if (inverse) {
        if (n_plan) return b200_set_err(B200FHE_EUNSUPPORTED, "nwt_backward_protected: fault injection is defined for the forward transform only");
        for (size_t i = start; i < start + L; i++)
            if (!c->has_inverse[i]) return b200_set_err(B200FHE_EINVAL, "nwt_backward_protected: no inverse tables for modulus %zu", i);
    }
    if (n_poly == 0 || L == 0) return 0;
    if (!d || !d_verdict) return b200_set_err(B200FHE_EINVAL, "nwt_protected: null argument");
    if (n_plan > B200FHE_MAX_INJECT || (n_plan && !plan)) return b200_set_err(B200FHE_EINVAL, "nwt_protected: bad injection plan");
    const size_t n_limbs = n_poly * L;
    CK(cudaSetDevice(c->device));
    cudaStream_t s = (cudaStream_t)stream;
    rc = ensure_abft(c, start, L, n_limbs, s);
    if (rc) return rc;
    // this call's slice of the rings
    const unsigned slot = c->abft_next++ % kAbftSlots;
    u64* d_cnt = (u64*)c->d_abft + (size_t)slot * 16 * c->abft_limbs;
    InjEntry* d_inj = c->d_inj + (size_t)slot * B200FHE_MAX_INJECT;
    if (c->abft_dirty & (1u << slot)) {         // a call on this slot failed midway: its counters may be non-zero
        CK(cudaMemsetAsync(d_cnt, 0, sizeof(uint64_t) * 16 * n_limbs, s));
        c->abft_dirty &= ~(1u << slot);
    }
    InjEntry h[B200FHE_MAX_INJECT];
    std::vector<b200fhe_inject> storage;        // kind 3 entries
    size_t n_bf = 0;                             // butterfly-product entries (kinds 1, 2)
    for (size_t i = 0; i < n_plan; i++) {
        if (plan[i].kind == 3) {               // storage fault between the passes: not a butterfly, handled below
            if (plan[i].limb >= n_limbs || plan[i].butterfly >= c->N || c->log_n < 12)
                return b200_set_err(B200FHE_EINVAL, "nwt_protected: storage fault %zu out of range (needs N >= 2^12)", i);
            storage.push_back(plan[i]);
            continue;
        }
        if (plan[i].limb >= n_limbs || plan[i].stage >= (uint32_t)c->log_n || plan[i].butterfly >= c->N / 2 || plan[i].kind < 1 || plan[i].kind > 2)
            return b200_set_err(B200FHE_EINVAL, "nwt_protected: injection %zu out of range", i);
        const uint64_t gap = c->N >> (plan[i].stage + 1);
        InjEntry* hp = h + n_bf++;
        hp->limb = plan[i].limb; hp->stage = plan[i].stage;
        hp->xa = (u32)(2 * (plan[i].butterfly / gap) * gap + plan[i].butterfly % gap);
        hp->kind = plan[i].kind; hp->payload = plan[i].payload;
    }
    if (n_bf) {
        CK(cudaMemcpyAsync(d_inj, h, sizeof(InjEntry) * n_bf, cudaMemcpyHostToDevice, s));
        CK(cudaStreamSynchronize(s));           // h[] is a stack buffer
    }
    c->abft_dirty |= (1u << slot);              // cleared once finalize (which zeroes the counters) is enqueued
    bool all_big = true, all_near = true;
    for (size_t i = start; i < start + L; i++) {
        all_big = all_big && c->h_mods[i].big;
        all_near = all_near && c->h_mods[i].near;
    }
    const int big = (all_near && !getenv("B200FHE_NO_NEAR")) ? 2 : all_big ? 1 : 0;
    PassArgs a;
    a.mods = c->d_mods;
    a.tw = inverse ? c->d_itw : c->d_tw;
    a.rtw = inverse ? c->d_irtw : c->d_rtw;
    a.N = (u32)c->N;
    a.L = (u32)L;
    a.start_mod = (u32)start;
    a.omega = (const u64*)c->d_omega;
    a.abft = d_cnt;
    a.inj = d_inj;
    a.n_inj = (u32)n_bf;
    size_t step = chunk_polys(c, L);
    if (step * L > 32768) step = 32768 / L ? 32768 / L : 1;
    for (size_t p0 = 0; p0 < n_poly; p0 += step) {
        const size_t np = n_poly - p0 < step ? n_poly - p0 : step;
        a.data = (u64*)d + p0 * L * c->N;
        a.limb0 = (u32)(p0 * L);
        if (!storage.empty()) {
            // pass 1, the stored intermediate words XOR-ed in HBM (the 2D transform's round trip: the SCF site of a GPU), pass 2
            rc = launch_ntt_protected(big, a, c->log_n, (u32)(np * L), s, 1);
            for (const b200fhe_inject& f : storage)
                if (!rc && f.limb >= p0 * L && f.limb < (p0 + np) * L) {
                    k_xor_word<<<1, 1, 0, s>>>((u64*)d + (size_t)f.limb * c->N + f.butterfly, f.payload);
                    g_launches += 1;
                }
            if (!rc) rc = launch_ntt_protected(big, a, c->log_n, (u32)(np * L), s, 2);
        } else
        rc = inverse ? launch_ntt_protected_inverse(big, a, c->log_n, (u32)(np * L), s) : launch_ntt_protected(big, a, c->log_n, (u32)(np * L), s);
        if (rc) return b200_set_err(B200FHE_EUNSUPPORTED, "nwt_protected: log_n=%d unsupported (rc=%d)", c->log_n, rc);
        g_launches += c->log_n <= 11 ? 1 : 2;
    }
    launch_abft_finalize(d_cnt, c->d_mods, (u32)L, (u32)start, (u32)n_limbs, d_verdict, (u64*)d_sums, s);
    g_launches += 1;
    CK(cudaGetLastError());
    c->abft_dirty &= ~(1u << slot);
    return 0;
}

extern "C" int b200fhe_nwt_forward_protected(b200fhe_ctx* c, uint64_t* d, size_t n_poly, size_t L, size_t start, uint32_t* d_verdict,
                                             uint64_t* d_sums, const b200fhe_inject* plan, size_t n_plan, void* stream) {
    return nwt_protected_impl(false, c, d, n_poly, L, start, d_verdict, d_sums, plan, n_plan, stream);
}
extern "C" int b200fhe_nwt_backward_protected(b200fhe_ctx* c, uint64_t* d, size_t n_poly, size_t L, size_t start, uint32_t* d_verdict,
                                              uint64_t* d_sums, void* stream) {
    return nwt_protected_impl(true, c, d, n_poly, L, start, d_verdict, d_sums, nullptr, 0, stream);
}

static int nwt_protected_host_impl(bool inverse, b200fhe_ctx* c, uint64_t* h, size_t n_poly, size_t L, size_t start, uint32_t* h_verdict,
                                   uint64_t* h_sums, const b200fhe_inject* plan, size_t n_plan) {
    int rc = check_limbs(c, L, start, "nwt_protected_host");
    if (rc) return rc;
    if (n_poly == 0 || L == 0) return 0;
    if (!h || !h_verdict) return b200_set_err(B200FHE_EINVAL, "nwt_protected_host: null argument");
    CK(cudaSetDevice(c->device));
    const size_t words = n_poly * L * c->N, n_limbs = n_poly * L;
    // staging: data | sums [n_limbs][2] | verdict (u32, padded to words)
    const size_t extra = 2 * n_limbs + (n_limbs + 1) / 2;
    rc = b200_ensure_stage(c, words + extra, false);
    if (rc) return rc;
    uint64_t* d_sums = c->d_stage + words;
    uint32_t* d_verdict = reinterpret_cast<uint32_t*>(d_sums + 2 * n_limbs);
    cudaStream_t s = c->own_stream;
    CK(cudaMemcpyAsync(c->d_stage, h, words * sizeof(uint64_t), cudaMemcpyHostToDevice, s));
    rc = nwt_protected_impl(inverse, c, c->d_stage, n_poly, L, start, d_verdict, d_sums, plan, n_plan, s);
    if (rc) return rc;
    CK(cudaMemcpyAsync(h, c->d_stage, words * sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(h_verdict, d_verdict, n_limbs * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    if (h_sums) CK(cudaMemcpyAsync(h_sums, d_sums, 2 * n_limbs * sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return 0;
}

extern "C" int b200fhe_nwt_forward_protected_host(b200fhe_ctx* c, uint64_t* h, size_t n_poly, size_t L, size_t start, uint32_t* h_verdict,
                                                  uint64_t* h_sums, const b200fhe_inject* plan, size_t n_plan) {
    return nwt_protected_host_impl(false, c, h, n_poly, L, start, h_verdict, h_sums, plan, n_plan);
}
extern "C" int b200fhe_nwt_backward_protected_host(b200fhe_ctx* c, uint64_t* h, size_t n_poly, size_t L, size_t start, uint32_t* h_verdict,
                                                   uint64_t* h_sums) {
    return nwt_protected_host_impl(true, c, h, n_poly, L, start, h_verdict, h_sums, nullptr, 0);
}

// ---------------------------------------------------------------- pmult
static int check_limbs(const b200fhe_ctx* c, size_t L, size_t start, const char* what) {
    if (!c) return b200_set_err(B200FHE_EINVAL, "%s: null ctx", what);
    if (start + L > (size_t)c->L) return b200_set_err(B200FHE_EINVAL, "%s: start_mod %zu + L %zu exceeds context (%d)", what, start, L, c->L);
    for (size_t i = start; i < start + L; i++)
        if (!c->moduli[i]) return b200_set_err(B200FHE_EINVAL, "%s: modulus %zu not set", what, i);
    return 0;
}

extern "C" int b200fhe_pmult_accum(b200fhe_ctx* c, const uint64_t* const* ct, const uint64_t* const* pt, size_t T, uint64_t* d_out,
                                   size_t P, size_t L, size_t start, int accumulate, void* stream) {
    int rc = check_limbs(c, L, start, "pmult");
    if (rc) return rc;
    if (P == 0 || L == 0) return 0;
    if (!d_out || (T && (!ct || !pt))) return b200_set_err(B200FHE_EINVAL, "pmult: null argument");
    for (size_t k = 0; k < T; k++)
        if (!ct[k] || !pt[k]) return b200_set_err(B200FHE_EINVAL, "pmult: null term %zu", k);
    CK(cudaSetDevice(c->device));
    unsigned long long nl_ = 0;
    rc = launch_pmult((const u64* const*)ct, (const u64* const*)pt, T, (u64*)d_out, c->d_mods, (u32)c->N, (u32)P, (u32)L, (u32)start,
                      accumulate, (cudaStream_t)stream, &nl_);
    g_launches += nl_;
    if (rc) return b200_set_err(B200FHE_EUNSUPPORTED, "pmult: N=%zu unsupported", c->N);
    CK(cudaGetLastError());
    return 0;
}

extern "C" int b200fhe_pmult_accum_host(b200fhe_ctx* c, const uint64_t* h_ct, const uint64_t* h_pt, size_t T, uint64_t* h_out, size_t P,
                                        size_t L, size_t start) {
    int rc = check_limbs(c, L, start, "pmult_host");
    if (rc) return rc;
    if (P == 0 || L == 0) return 0;
    if (!h_out || (T && (!h_ct || !h_pt))) return b200_set_err(B200FHE_EINVAL, "pmult_host: null argument");
    CK(cudaSetDevice(c->device));
    const size_t ptw = L * c->N, ctw = P * ptw;
    // chunk of terms per buffer: about 256 MiB of input, two buffers in flight
    size_t tc = ((size_t)256 << 20) / ((ctw + ptw) * sizeof(uint64_t));
    if (tc < 1) tc = 1;
    if (tc > 64) tc = 64;
    if (T && tc > T) tc = T;
    const size_t bufw = tc * (ctw + ptw);
    rc = b200_ensure_stage(c, 2 * bufw + ctw, false);
    if (rc) return rc;
    if (!c->s_h2d) {
        CK(cudaStreamCreateWithFlags(&c->s_h2d, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&c->s_d2h, cudaStreamNonBlocking));
    }
    if (c->ev.size() < 4) {
        size_t old = c->ev.size();
        c->ev.resize(4);
        for (size_t i = old; i < 4; i++) CK(cudaEventCreateWithFlags(&c->ev[i], cudaEventDisableTiming));
    }
    uint64_t* d_out = c->d_stage + 2 * bufw;
    cudaStream_t sk = c->own_stream;
    std::vector<const uint64_t*> pc(tc), pp(tc);
    if (T == 0) {
        rc = b200fhe_pmult_accum(c, nullptr, nullptr, 0, d_out, P, L, start, 0, sk);
        if (rc) return rc;
    }
    size_t chunk = 0;
    for (size_t k0 = 0; k0 < T; k0 += tc, chunk++) {
        const size_t n = T - k0 < tc ? T - k0 : tc;
        uint64_t* buf = c->d_stage + (chunk & 1) * bufw;
        if (chunk >= 2) CK(cudaStreamWaitEvent(c->s_h2d, c->ev[2 + (chunk & 1)], 0));   // kernel that last read this buffer
        CK(cudaMemcpyAsync(buf, h_ct + k0 * ctw, n * ctw * sizeof(uint64_t), cudaMemcpyHostToDevice, c->s_h2d));
        CK(cudaMemcpyAsync(buf + tc * ctw, h_pt + k0 * ptw, n * ptw * sizeof(uint64_t), cudaMemcpyHostToDevice, c->s_h2d));
        CK(cudaEventRecord(c->ev[chunk & 1], c->s_h2d));
        CK(cudaStreamWaitEvent(sk, c->ev[chunk & 1], 0));
        for (size_t k = 0; k < n; k++) { pc[k] = buf + k * ctw; pp[k] = buf + tc * ctw + k * ptw; }
        rc = b200fhe_pmult_accum(c, pc.data(), pp.data(), n, d_out, P, L, start, k0 != 0, sk);
        if (rc) return rc;
        CK(cudaEventRecord(c->ev[2 + (chunk & 1)], sk));
    }
    CK(cudaMemcpyAsync(h_out, d_out, ctw * sizeof(uint64_t), cudaMemcpyDeviceToHost, sk));
    CK(cudaStreamSynchronize(sk));
    return 0;
}

// ---------------------------------------------------------------- base conversion
extern "C" int b200fhe_bconv_destroy(b200fhe_bconv* p) {
    if (!p) return 0;
    if (p->ctx) cudaSetDevice(p->ctx->device);
    if (p->d_src) cudaFree(p->d_src);
    if (p->d_dst) cudaFree(p->d_dst);
    if (p->d_inv) cudaFree(p->d_inv);
    if (p->d_hat) cudaFree(p->d_hat);
    if (p->d_btab) cudaFree(p->d_btab);
    if (p->d_btab5) cudaFree(p->d_btab5);
    if (p->d_hatw) cudaFree(p->d_hatw);
    if (p->d_ePw) cudaFree(p->d_ePw);
    if (p->d_ePq) cudaFree(p->d_ePq);
    if (p->d_cnt) cudaFree(p->d_cnt);
    if (p->d_y) cudaFree(p->d_y);
    if (p->d_io) cudaFree(p->d_io);
    delete p;
    return 0;
}

extern "C" int b200fhe_bconv_create(b200fhe_ctx* c, const int* src_idx, int m, const int* dst_idx, int K, b200fhe_bconv** out) {
    return b200fhe_bconv_create_scaled(c, src_idx, m, dst_idx, K, 1, out);
}

// t > 1: the conversion returns t * [x t^-1]_P (+ e t P): congruent to x modulo P and to 0 modulo t -- what BGV's ModDown /
// mod-switch subtract so that the division by P leaves the plaintext (mod t) intact.  t = 1: plain HPS conversion.
extern "C" int b200fhe_bconv_create_scaled(b200fhe_ctx* c, const int* src_idx, int m, const int* dst_idx, int K, uint64_t t,
                                           b200fhe_bconv** out) {
    if (!c || !src_idx || !dst_idx || !out || m < 1 || K < 1 || m > 64 || t < 1) return b200_set_err(B200FHE_EINVAL, "bconv_create: bad arguments");
    std::vector<u32> src(m), dst(K);
    for (int j = 0; j < m; j++) {
        if (src_idx[j] < 0 || src_idx[j] >= c->L || !c->moduli[src_idx[j]]) return b200_set_err(B200FHE_EINVAL, "bconv_create: bad source index %d", src_idx[j]);
        for (int i = 0; i < j; i++)
            if (src_idx[i] == src_idx[j]) return b200_set_err(B200FHE_EINVAL, "bconv_create: repeated source modulus");
        src[j] = (u32)src_idx[j];
    }
    for (int k = 0; k < K; k++) {
        if (dst_idx[k] < 0 || dst_idx[k] >= c->L || !c->moduli[dst_idx[k]]) return b200_set_err(B200FHE_EINVAL, "bconv_create: bad target index %d", dst_idx[k]);
        dst[k] = (u32)dst_idx[k];
    }
    std::vector<ulonglong2> inv(m);
    std::vector<uint64_t> hat((size_t)K * m), raw((size_t)K * m);
    for (int j = 0; j < m; j++) {
        const uint64_t pj = c->moduli[src[j]];
        uint64_t h = 1;
        for (int i = 0; i < m; i++)
            if (i != j) h = nt::mulmod(h, c->moduli[src[i]] % pj, pj);
        if (t % pj == 0) return b200_set_err(B200FHE_EINVAL, "bconv_create: scale %llu is not invertible modulo source modulus %d", (unsigned long long)t, src_idx[j]);
        const uint64_t iv = nt::invmod(nt::mulmod(h, t % pj, pj), pj);
        inv[j] = make_ulonglong2(iv, nt::shoup(iv, pj));
        for (int k = 0; k < K; k++) {
            const uint64_t qk = c->moduli[dst[k]];
            uint64_t g = 1;
            for (int i = 0; i < m; i++)
                if (i != j) g = nt::mulmod(g, c->moduli[src[i]] % qk, qk);
            g = nt::mulmod(g, t % qk, qk);
            hat[(size_t)k * m + j] = (g & 0x3fffffffull) | ((g >> 30) << 32);
            raw[(size_t)k * m + j] = g;
        }
    }
    CK(cudaSetDevice(c->device));
    b200fhe_bconv* p = new b200fhe_bconv();
    p->ctx = c; p->m = m; p->K = K;
    cudaError_t e = cudaMalloc(&p->d_src, sizeof(u32) * m);
    if (e == cudaSuccess) e = cudaMalloc(&p->d_dst, sizeof(u32) * K);
    if (e == cudaSuccess) e = cudaMalloc(&p->d_inv, sizeof(ulonglong2) * m);
    if (e == cudaSuccess) e = cudaMalloc(&p->d_hat, sizeof(uint64_t) * K * m);
    if (e == cudaSuccess) e = cudaMalloc(&p->d_y, sizeof(uint64_t) * m * c->N);
    if (e == cudaSuccess) e = cudaMemcpy(p->d_src, src.data(), sizeof(u32) * m, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(p->d_dst, dst.data(), sizeof(u32) * K, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(p->d_inv, inv.data(), sizeof(ulonglong2) * m, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(p->d_hat, hat.data(), sizeof(uint64_t) * K * m, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && m <= 16 && !getenv("B200FHE_BCONV_NO_TC")) {
        // B fragments of mma.m16n8k32 (col-major 32 x 8, lane = 4 g' + t holds k = 4t..4t+3 (+16) of column g'), laid out
        // [n-tile u][k-step s][target][e = (g' & 1) * 4 + t]: column g' of n-tile u = digit 2u + (g' & 1) of target base + g'/2
        const int KS = (m + 3) / 4;
        std::vector<uint2> bt((size_t)8 * KS * K * 8);
        for (int u = 0; u < 8; u++)
            for (int s = 0; s < KS; s++)
                for (int k = 0; k < K; k++)
                    for (int el = 0; el < 8; el++) {
                        const int digit = 2 * u + (el >> 2), t = el & 3;
                        uint32_t w[2] = {0, 0};
                        for (int half = 0; half < 2; half++)
                            for (int cb = 0; cb < 4; cb++) {
                                const int kidx = 32 * s + 16 * half + 4 * t + cb, j = kidx / 8, a = kidx % 8, sh = digit - a;
                                uint32_t byte = 0;
                                if (j < m && sh >= 0 && sh < 8) byte = (uint32_t)((raw[(size_t)k * m + j] >> (8 * sh)) & 0xff);
                                w[half] |= byte << (8 * cb);
                            }
                        bt[(((size_t)u * KS + s) * K + k) * 8 + el] = make_uint2(w[0], w[1]);
                    }
        e = cudaMalloc(&p->d_btab, sizeof(uint2) * bt.size());
        if (e == cudaSuccess) e = cudaMemcpy(p->d_btab, bt.data(), sizeof(uint2) * bt.size(), cudaMemcpyHostToDevice);
        // the same constants as rows of the tcgen05 B operand: [target][16-byte K chunk kc][digit d] -> bytes (j, a) = (2 kc + b / 8, b % 8)
        std::vector<uint8_t> b5((size_t)K * 8 * 16 * 16, 0);
        for (int k = 0; k < K; k++)
            for (int kc = 0; kc < 8; kc++)
                for (int d = 0; d < 16; d++)
                    for (int b = 0; b < 16; b++) {
                        const int j = 2 * kc + b / 8, a = b % 8, sh = d - a;
                        if (j < m && sh >= 0 && sh < 8)
                            b5[(((size_t)k * 8 + kc) * 16 + d) * 16 + b] = (uint8_t)((raw[(size_t)k * m + j] >> (8 * sh)) & 0xff);
                    }
        if (e == cudaSuccess) e = cudaMalloc(&p->d_btab5, b5.size());
        if (e == cudaSuccess) e = cudaMemcpy(p->d_btab5, b5.data(), b5.size(), cudaMemcpyHostToDevice);
    }
    if (e != cudaSuccess) {
        b200fhe_bconv_destroy(p);
        return b200_set_err(B200FHE_ECUDA, "bconv_create: %s", cudaGetErrorString(e));
    }
    *out = p;
    return 0;
}

// Exact-CRT mode (SURVEY 8(b) `mode /*HPS|EXACT*/`; motivation/baseConv.py:67-83 base_conv_fixed): the plan additionally holds
// P/p_j and the multiples e P as multi-word integers and e P mod q_k; b200fhe_bconv_modup[_host] then return
// (x mod P) mod q_k exactly instead of (x + e P) mod q_k.
extern "C" int b200fhe_bconv_create_mode(b200fhe_ctx* c, const int* src_idx, int m, const int* dst_idx, int K, int mode, b200fhe_bconv** out) {
    if (mode != B200FHE_BCONV_HPS && mode != B200FHE_BCONV_EXACT) return b200_set_err(B200FHE_EINVAL, "bconv_create_mode: mode must be HPS (0) or EXACT (1)");
    int rc = b200fhe_bconv_create_scaled(c, src_idx, m, dst_idx, K, 1, out);
    if (rc || mode == B200FHE_BCONV_HPS) return rc;
    b200fhe_bconv* p = *out;
    // little-endian multi-word integers, m words each
    auto mul_small = [&](std::vector<uint64_t>& a, uint64_t f) {
        uint64_t carry = 0;
        for (auto& w : a) { nt::u128 t = (nt::u128)w * f + carry; w = (uint64_t)t; carry = (uint64_t)(t >> 64); }
    };
    auto mod_small = [&](const std::vector<uint64_t>& a, uint64_t q) {
        nt::u128 r = 0;
        for (size_t i = a.size(); i-- > 0;) r = ((r << 64) | a[i]) % q;
        return (uint64_t)r;
    };
    std::vector<uint64_t> hatw((size_t)m * m, 0), ePw((size_t)m * m, 0), ePq((size_t)K * m, 0);
    std::vector<uint64_t> P(m, 0);
    P[0] = 1;
    for (int j = 0; j < m; j++) mul_small(P, c->moduli[src_idx[j]]);
    for (int j = 0; j < m; j++) {
        std::vector<uint64_t> h(m, 0);
        h[0] = 1;
        for (int i = 0; i < m; i++) if (i != j) mul_small(h, c->moduli[src_idx[i]]);
        std::copy(h.begin(), h.end(), hatw.begin() + (size_t)j * m);
    }
    for (int e = 1; e < m; e++) {
        std::vector<uint64_t> w = P;
        mul_small(w, (uint64_t)e);                   // e P < m 2^(60 m) fits m words
        std::copy(w.begin(), w.end(), ePw.begin() + (size_t)e * m);
        for (int k = 0; k < K; k++) ePq[(size_t)k * m + e] = mod_small(w, c->moduli[dst_idx[k]]);
    }
    cudaError_t e = cudaMalloc(&p->d_hatw, sizeof(uint64_t) * hatw.size());
    if (e == cudaSuccess) e = cudaMalloc(&p->d_ePw, sizeof(uint64_t) * ePw.size());
    if (e == cudaSuccess) e = cudaMalloc(&p->d_ePq, sizeof(uint64_t) * ePq.size());
    if (e == cudaSuccess) e = cudaMemcpy(p->d_hatw, hatw.data(), sizeof(uint64_t) * hatw.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(p->d_ePw, ePw.data(), sizeof(uint64_t) * ePw.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(p->d_ePq, ePq.data(), sizeof(uint64_t) * ePq.size(), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        b200fhe_bconv_destroy(p);
        *out = nullptr;
        return b200_set_err(B200FHE_ECUDA, "bconv_create_mode: %s", cudaGetErrorString(e));
    }
    p->mode = B200FHE_BCONV_EXACT;
    return 0;
}

extern "C" int b200fhe_bconv_scale(b200fhe_bconv* p, const uint64_t* d_x, uint64_t* d_y, size_t j0, size_t count, void* stream) {
    if (!p || (count && (!d_x || !d_y)) || j0 + count > (size_t)p->m) return b200_set_err(B200FHE_EINVAL, "bconv_scale: bad arguments");
    CK(cudaSetDevice(p->ctx->device));
    launch_bconv_scale((const u64*)d_x, (u64*)d_y, p->ctx->d_mods, p->d_src, p->d_inv, (u32)p->ctx->N, (u32)j0, (u32)count, (cudaStream_t)stream);
    if (count) g_launches++;
    CK(cudaGetLastError());
    return 0;
}

extern "C" int b200fhe_bconv_contract(b200fhe_bconv* p, const uint64_t* d_y, uint64_t* d_out, size_t k0, size_t kcount, void* stream) {
    if (!p || (kcount && (!d_y || !d_out)) || k0 + kcount > (size_t)p->K) return b200_set_err(B200FHE_EINVAL, "bconv_contract: bad arguments");
    CK(cudaSetDevice(p->ctx->device));
    int rc = -1;
    if (p->d_btab5)
        rc = launch_bconv_contract_tc5((const u64*)d_y, (u64*)d_out, p->ctx->d_mods, p->d_dst, p->d_btab5, (u32)p->ctx->N, (u32)p->m, (u32)p->K,
                                       (u32)k0, (u32)kcount, (cudaStream_t)stream);
    if (rc == 0) {
        if (kcount) g_launches += (kcount + 47) / 48;
        CK(cudaGetLastError());
        return 0;
    }
    if (p->d_btab)
        rc = launch_bconv_contract_tc((const u64*)d_y, (u64*)d_out, p->ctx->d_mods, p->d_dst, p->d_btab, (u32)p->ctx->N, (u32)p->m, (u32)p->K,
                                      (u32)k0, (u32)kcount, (cudaStream_t)stream);
    else
        rc = launch_bconv_contract((const u64*)d_y, (u64*)d_out, p->ctx->d_mods, p->d_dst, (const u64*)p->d_hat, (u32)p->ctx->N, (u32)p->m, (u32)p->K,
                                   (u32)k0, (u32)kcount, (cudaStream_t)stream);
    if (rc) return b200_set_err(B200FHE_EUNSUPPORTED, "bconv_contract: source base too large for shared memory (m=%d)", p->m);
    if (kcount) g_launches++;
    CK(cudaGetLastError());
    return 0;
}

// All K targets of the plan in one launch, written around a gap: target t lands in row t of d_out for t < skip_at and in row
// t + skip_rows from there on (the key-switch's raised polynomial keeps the digit's own limbs in that gap).  Returns 1 when the
// tcgen05 contraction does not serve the shape: the caller then issues the two range calls.
int b200_bconv_contract_skip(b200fhe_bconv* p, const uint64_t* d_y, uint64_t* d_out, size_t skip_at, size_t skip_rows, void* stream) {
    if (!p || !d_y || !d_out) return b200_set_err(B200FHE_EINVAL, "bconv_contract: bad arguments");
    if (!p->d_btab5) return 1;
    CK(cudaSetDevice(p->ctx->device));
    if (launch_bconv_contract_tc5_skip((const u64*)d_y, (u64*)d_out, p->ctx->d_mods, p->d_dst, p->d_btab5, (u32)p->ctx->N, (u32)p->m, (u32)p->K,
                                       (u32)skip_at, (u32)skip_rows, (cudaStream_t)stream))
        return 1;
    g_launches += ((size_t)p->K + 47) / 48;
    CK(cudaGetLastError());
    return 0;
}

extern "C" int b200fhe_bconv_modup(b200fhe_bconv* p, const uint64_t* d_in, uint64_t* d_out, void* stream) {
    if (!p) return b200_set_err(B200FHE_EINVAL, "bconv_modup: null plan");
    int rc = b200fhe_bconv_scale(p, d_in, (uint64_t*)p->d_y, 0, p->m, stream);
    if (rc) return rc;
    rc = b200fhe_bconv_contract(p, (const uint64_t*)p->d_y, d_out, 0, p->K, stream);
    if (rc || p->mode != B200FHE_BCONV_EXACT) return rc;
    if (launch_bconv_exact_fix(p->d_y, (u64*)d_out, (u32)p->ctx->N, (u32)p->m, (u32)p->K, p->d_hatw, p->d_ePw, p->d_ePq, p->ctx->d_mods, p->d_dst,
                               (cudaStream_t)stream))
        return b200_set_err(B200FHE_EUNSUPPORTED, "bconv_modup: exact mode supports at most 64 source moduli");
    g_launches++;
    CK(cudaGetLastError());
    return 0;
}

extern "C" int b200fhe_bconv_modup_host(b200fhe_bconv* p, const uint64_t* h_in, uint64_t* h_out) {
    if (!p || !h_in || !h_out) return b200_set_err(B200FHE_EINVAL, "bconv_modup_host: null argument");
    b200fhe_ctx* c = p->ctx;
    CK(cudaSetDevice(c->device));
    const size_t wi = (size_t)p->m * c->N, wo = (size_t)p->K * c->N;
    if (p->io_words < wi + wo) {
        if (p->d_io) cudaFree(p->d_io);
        p->d_io = nullptr; p->io_words = 0;
        CK(cudaMalloc(&p->d_io, sizeof(uint64_t) * (wi + wo)));
        p->io_words = wi + wo;
    }
    cudaStream_t s = c->own_stream;
    CK(cudaMemcpyAsync(p->d_io, h_in, sizeof(uint64_t) * wi, cudaMemcpyHostToDevice, s));
    int rc = b200fhe_bconv_modup(p, (const uint64_t*)p->d_io, (uint64_t*)(p->d_io + wi), s);
    if (rc) return rc;
    CK(cudaMemcpyAsync(h_out, p->d_io + wi, sizeof(uint64_t) * wo, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return 0;
}

int b200_bconv_modup_protected_split(b200fhe_bconv* p, const uint64_t* d_in, uint64_t* d_y, uint64_t* d_out, uint64_t* d_out_hi, size_t k_split,
                                     uint32_t* d_verdict, uint64_t* d_sums, const b200fhe_bconv_inject* inject, size_t n_inject, void* stream);

extern "C" int b200fhe_bconv_modup_protected(b200fhe_bconv* p, const uint64_t* d_in, uint64_t* d_out, uint32_t* d_verdict, uint64_t* d_sums,
                                             const b200fhe_bconv_inject* inject, size_t n_inject, void* stream) {
    return b200_bconv_modup_protected_split(p, d_in, p ? (uint64_t*)p->d_y : nullptr, d_out, nullptr, 0, d_verdict, d_sums, inject, n_inject, stream);
}

// shared with keyswitch.cu: scaled digit into d_y, targets [0, k_split) into d_out and the rest into d_out_hi (or all into d_out)
int b200_bconv_modup_protected_split(b200fhe_bconv* p, const uint64_t* d_in, uint64_t* d_y, uint64_t* d_out, uint64_t* d_out_hi, size_t k_split,
                                     uint32_t* d_verdict, uint64_t* d_sums, const b200fhe_bconv_inject* inject, size_t n_inject, void* stream) {
    if (!p || !d_in || !d_y || !d_out || !d_verdict) return b200_set_err(B200FHE_EINVAL, "bconv_modup_protected: null argument");
    if (n_inject > B200FHE_BCONV_MAX_INJECT || (n_inject && !inject)) return b200_set_err(B200FHE_EINVAL, "bconv_modup_protected: bad injection plan");
    if (!p->d_btab) return b200_set_err(B200FHE_EUNSUPPORTED, "bconv_modup_protected: needs the tensor-core plan (at most 16 source moduli)");
    CK(cudaSetDevice(p->ctx->device));
    const size_t n_cnt = 6 * (size_t)p->m + 2 * (size_t)p->K;
    if (!p->d_cnt) {
        CK(cudaMalloc(&p->d_cnt, sizeof(uint64_t) * 2 * n_cnt));
        CK(cudaMemsetAsync(p->d_cnt, 0, sizeof(uint64_t) * 2 * n_cnt, (cudaStream_t)stream));
        p->cnt_set = 0;
    }
    u64* cur = p->d_cnt + (size_t)p->cnt_set * n_cnt;
    u64* nxt = p->d_cnt + (size_t)(p->cnt_set ^ 1) * n_cnt;
    p->cnt_set ^= 1;
    static_assert(sizeof(b200fhe_bconv_inject) == 24, "layout shared with BconvInj");
    unsigned long long nl2_ = 0;
    int rc = launch_bconv_modup_protected((const u64*)d_in, (u64*)d_y, (u64*)d_out, p->ctx->d_mods, p->d_src, p->d_dst, p->d_inv, p->d_hat, p->d_btab,
                                          (u32)p->ctx->N, (u32)p->m, (u32)p->K, cur, nxt, d_verdict, (u64*)d_sums, inject, (u32)n_inject,
                                          (cudaStream_t)stream, &nl2_, (u64*)d_out_hi, (u32)k_split, p->d_btab5);
    g_launches += nl2_;
    if (rc) return b200_set_err(B200FHE_EUNSUPPORTED, "bconv_modup_protected: unsupported plan");
    CK(cudaGetLastError());
    return 0;
}

extern "C" int b200fhe_bconv_modup_protected_host(b200fhe_bconv* p, const uint64_t* h_in, uint64_t* h_out, uint32_t* h_verdict,
                                                  uint64_t* h_sums, const b200fhe_bconv_inject* inject, size_t n_inject) {
    if (!p || !h_in || !h_out || !h_verdict) return b200_set_err(B200FHE_EINVAL, "bconv_modup_protected_host: null argument");
    b200fhe_ctx* c = p->ctx;
    CK(cudaSetDevice(c->device));
    const size_t wi = (size_t)p->m * c->N, wo = (size_t)p->K * c->N, nchk = (size_t)p->m + p->K;
    const size_t need = wi + wo + 2 * nchk + (nchk + 1) / 2;          // data | sums | verdict (u32)
    if (p->io_words < need) {
        if (p->d_io) cudaFree(p->d_io);
        p->d_io = nullptr; p->io_words = 0;
        CK(cudaMalloc(&p->d_io, sizeof(uint64_t) * need));
        p->io_words = need;
    }
    cudaStream_t s = c->own_stream;
    u64* d_sums = p->d_io + wi + wo;
    uint32_t* d_ver = reinterpret_cast<uint32_t*>(d_sums + 2 * nchk);
    CK(cudaMemcpyAsync(p->d_io, h_in, sizeof(uint64_t) * wi, cudaMemcpyHostToDevice, s));
    int rc = b200fhe_bconv_modup_protected(p, (const uint64_t*)p->d_io, (uint64_t*)(p->d_io + wi), d_ver, (uint64_t*)d_sums, inject, n_inject, s);
    if (rc) return rc;
    CK(cudaMemcpyAsync(h_out, p->d_io + wi, sizeof(uint64_t) * wo, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(h_verdict, d_ver, sizeof(uint32_t) * nchk, cudaMemcpyDeviceToHost, s));
    if (h_sums) CK(cudaMemcpyAsync(h_sums, d_sums, sizeof(uint64_t) * 2 * nchk, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return 0;
}

// ---------------------------------------------------------------- fault injection
__global__ void k_flip_bit(u64* data, size_t idx, unsigned bit) { data[idx] ^= (1ull << bit); }

extern "C" int b200fhe_flip_bit(uint64_t* d, size_t idx, unsigned bit, void* stream) {
    if (!d || bit > 63) return b200_set_err(B200FHE_EINVAL, "flip_bit: bad args");
    k_flip_bit<<<1, 1, 0, (cudaStream_t)stream>>>((u64*)d, idx, bit);
    CK(cudaGetLastError());
    return 0;
}
