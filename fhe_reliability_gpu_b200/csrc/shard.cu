// Limb-sharded key-switch ModUp over NCCL / NVLink (config C5, SURVEY 8(e) and the b200fhe_shard_* / b200fhe_modup_sharded
// entries of SURVEY 8(b)).  The reference has no multi-GPU code; this is the step Phantom's relinearize_inplace /
// rotate_inplace run on one device (reliability_test/dotprod_test.cu:114,146; NTT 48 % + base conversion 42 % of a switch at
// dnum = 3, profile_framewk/draw_dnum.py:5-9) split over `world` GPUs, one process per GPU.
//
// Ownership: limb g of Q u P (global index, q_0..q_{n_q-1}, p_0..p_{n_p-1}) lives on rank g % world.  A rank holds its
// limbs in ascending order; its private context lists them FIRST, so row i of a rank's buffers is modulus i of that
// context and every batched launch covers a contiguous modulus run.  One call:
//   1. x = INTT(own Q limbs) with the conversions' scale step folded into the transform's final constant  (1 launch pair)
//   2. per digit d, on the communication stream: all-gather of the digit's scaled limbs, rank-major         (THE collective)
//   3. per digit d, as soon as its gather has landed: contraction into the owned target limbs + forward NTT; the digit's
//      own limbs pass through.  Gather d+1 overlaps contraction / NTT d (events, no host synchronisation), and the whole
//      call is capturable in a CUDA graph (fork / join of the two streams inside the capture).
// Equal per-rank counts use ncclAllGather; ragged digits (alpha % world != 0) a group of ncclBroadcast, so the gathered
// digit is always the compact [alpha][N] source matrix of the conversion, no padding rows.
// NCCL is resolved at run time (dlopen of libnccl.so.2: the copy torch already mapped, else the system one), so the
// library loads and every single-GPU entry works on a box without NCCL; world = 1 needs no communicator at all.
#include "../../include/b200fhe.h"
#include "ctx.hpp"
#include "numth.hpp"
#include <dlfcn.h>
#include <cstdlib>
#include <cstring>
#include <vector>

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

// ---- the slice of the NCCL API this file uses, bound with dlsym (types as in nccl.h 2.x) ------------------------------
namespace {
typedef struct { char internal[128]; } NcclUniqueId;
struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(NcclUniqueId*) = nullptr;
    int (*CommInitRank)(void**, int, NcclUniqueId, int) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    int (*CommCount)(void*, int*) = nullptr;
    int (*CommUserRank)(void*, int*) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int*) = nullptr;
};
constexpr int kNcclUint64 = 5;   // ncclUint64 in ncclDataType_t

NcclApi* nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api.handle ? &api : nullptr;
    tried = true;
    const char* names[] = {getenv("B200FHE_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        if (!n || !*n) continue;
        api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (api.handle) break;
    }
    if (!api.handle) return nullptr;
#define BIND(field, sym)                                                        \
    *(void**)(&api.field) = dlsym(api.handle, sym);                             \
    if (!api.field) { dlclose(api.handle); api.handle = nullptr; return nullptr; }
    BIND(GetUniqueId, "ncclGetUniqueId")
    BIND(CommInitRank, "ncclCommInitRank")
    BIND(CommDestroy, "ncclCommDestroy")
    BIND(CommCount, "ncclCommCount")
    BIND(CommUserRank, "ncclCommUserRank")
    BIND(AllGather, "ncclAllGather")
    BIND(Broadcast, "ncclBroadcast")
    BIND(GroupStart, "ncclGroupStart")
    BIND(GroupEnd, "ncclGroupEnd")
    BIND(GetErrorString, "ncclGetErrorString")
    BIND(GetVersion, "ncclGetVersion")
#undef BIND
    return &api;
}
}  // namespace

#define NC(api, call)                                                                                          \
    do {                                                                                                       \
        int r_ = (call);                                                                                       \
        if (r_ != 0) return b200_set_err(B200FHE_ECUDA, "%s failed: %s", #call, (api)->GetErrorString(r_));    \
    } while (0)

struct b200fhe_shard {
    void* comm = nullptr;        // ncclComm_t (NULL when world == 1)
    bool own_comm = false;
    int rank = 0, world = 1, device = 0;
    cudaStream_t comm_stream = nullptr;
};

struct b200fhe_modup_plan {
    b200fhe_shard* sh = nullptr;
    b200fhe_ctx* ctx = nullptr;                    // owned limbs first (ascending global index), then the others
    int n_q = 0, n_p = 0, dnum = 0, alpha = 0, total = 0;
    std::vector<int> own_q, own_all;               // global indices of the owned limbs of Q / of Q u P, ascending
    std::vector<std::vector<int>> counts;          // [dnum][world] digit limbs per rank
    std::vector<int> digit_row0, digit_rows;       // first owned row of digit d among own_q, and how many
    std::vector<b200fhe_bconv*> conv;              // per digit: source = the digit in gather order, targets = owned non-digit limbs
    std::vector<int> below;                        // per digit: owned target limbs that precede the digit's own rows
    B200Mod* d_mods_scaled = nullptr;              // the context's constants with N^-1 (D_d / q_j)^-1 for the owned Q limbs
    uint64_t* d_x = nullptr;                       // [n_own_q][N] scaled owned limbs (the send buffers are slices of it)
    uint64_t* d_yfull = nullptr;                   // [dnum][alpha][N] gathered digits
    cudaEvent_t ev_scaled = nullptr;
    std::vector<cudaEvent_t> ev_gathered;
    // the converted rows of every digit as one scatter list for the forward transform ((d n_own + r, r) for the rows r outside the
    // digit): with the limbs spread over the ranks a digit's launch is a fraction of a wave (6 limbs per digit on 8 GPUs = a third
    // of the column pass's CTA slots), so the transforms of all digits go out together after the last contraction
    std::vector<uint2> up_map;
    uint2* d_up_map = nullptr;
};

// ---- pure host layout (also what the CPU tests compare with modup.ShardPlan) ----------------------------------------
extern "C" int b200fhe_modup_sharded_layout(int n_q, int n_p, int dnum, int world, int rank, int* n_own_q, int* n_own_total,
                                            int* own_limbs, int* source_order, int* counts) {
    if (n_q < 1 || n_p < 0 || dnum < 1 || n_q % dnum || world < 1 || rank < 0 || rank >= world)
        return b200_set_err(B200FHE_EINVAL, "modup_sharded_layout: bad arguments");
    const int alpha = n_q / dnum, total = n_q + n_p;
    int nq = 0, nt_ = 0;
    for (int g = 0; g < total; g++)
        if (g % world == rank) {
            if (own_limbs) own_limbs[nt_] = g;
            nt_++;
            if (g < n_q) nq++;
        }
    if (n_own_q) *n_own_q = nq;
    if (n_own_total) *n_own_total = nt_;
    for (int d = 0; d < dnum; d++) {
        int o = 0;
        for (int r = 0; r < world; r++) {
            int c = 0;
            for (int g = d * alpha; g < (d + 1) * alpha; g++)
                if (g % world == r) {
                    if (source_order) source_order[d * alpha + o] = g;
                    o++; c++;
                }
            if (counts) counts[d * world + r] = c;
        }
    }
    return 0;
}

// ---- communicator -----------------------------------------------------------------------------------------------------
extern "C" int b200fhe_shard_unique_id(uint8_t id[128]) {
    NcclApi* api = nccl_api();
    if (!api) return b200_set_err(B200FHE_EUNSUPPORTED, "shard_unique_id: libnccl.so.2 not found (set B200FHE_NCCL_LIB)");
    NcclUniqueId u;
    NC(api, api->GetUniqueId(&u));
    memcpy(id, u.internal, 128);
    return 0;
}

static int shard_finish(b200fhe_shard* s, b200fhe_shard** out) {
    cudaError_t e = cudaSetDevice(s->device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s->comm_stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        delete s;
        return b200_set_err(B200FHE_ECUDA, "shard_init: %s", cudaGetErrorString(e));
    }
    *out = s;
    return 0;
}

// Adopts the caller's communicator (a C++ host that already runs NCCL): `nccl_comm` is an ncclComm_t.  world = 1 may pass NULL.
extern "C" int b200fhe_shard_init(void* nccl_comm, int rank, int world, b200fhe_shard** out) {
    if (!out || world < 1 || rank < 0 || rank >= world) return b200_set_err(B200FHE_EINVAL, "shard_init: bad arguments");
    b200fhe_shard* s = new b200fhe_shard();
    s->comm = nccl_comm; s->rank = rank; s->world = world;
    if (cudaGetDevice(&s->device) != cudaSuccess) { delete s; return b200_set_err(B200FHE_ECUDA, "shard_init: no CUDA device"); }
    if (world > 1 && nccl_comm) {       // NULL with world > 1: caller-provided transport, only the scale / raise halves work
        NcclApi* api = nccl_api();
        if (!api) { delete s; return b200_set_err(B200FHE_EUNSUPPORTED, "shard_init: libnccl.so.2 not found"); }
        int n = 0, r = -1;
        if (api->CommCount(nccl_comm, &n) || api->CommUserRank(nccl_comm, &r) || n != world || r != rank) {
            delete s;
            return b200_set_err(B200FHE_EINVAL, "shard_init: communicator is rank %d of %d, not %d of %d", r, n, rank, world);
        }
    }
    return shard_finish(s, out);
}

// Creates the communicator itself from an id made by b200fhe_shard_unique_id on one rank and handed to all
// (ncclCommInitRank; collective: every rank of the world must call it).
extern "C" int b200fhe_shard_init_id(const uint8_t id[128], int rank, int world, int device, b200fhe_shard** out) {
    if (!out || world < 1 || rank < 0 || rank >= world) return b200_set_err(B200FHE_EINVAL, "shard_init_id: bad arguments");
    b200fhe_shard* s = new b200fhe_shard();
    s->rank = rank; s->world = world; s->device = device;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) { delete s; return b200_set_err(B200FHE_ECUDA, "shard_init_id: %s", cudaGetErrorString(e)); }
    if (world > 1) {
        NcclApi* api = nccl_api();
        if (!api || !id) { delete s; return b200_set_err(B200FHE_EUNSUPPORTED, "shard_init_id: libnccl.so.2 not found or null id"); }
        NcclUniqueId u;
        memcpy(u.internal, id, 128);
        int r = api->CommInitRank(&s->comm, world, u, rank);
        if (r) { delete s; return b200_set_err(B200FHE_ECUDA, "ncclCommInitRank: %s", api->GetErrorString(r)); }
        s->own_comm = true;
    }
    return shard_finish(s, out);
}

extern "C" int b200fhe_shard_destroy(b200fhe_shard* s) {
    if (!s) return 0;
    cudaSetDevice(s->device);
    if (s->comm_stream) { cudaStreamSynchronize(s->comm_stream); cudaStreamDestroy(s->comm_stream); }
    if (s->own_comm && s->comm) { NcclApi* api = nccl_api(); if (api) api->CommDestroy(s->comm); }
    delete s;
    return 0;
}

extern "C" int b200fhe_shard_info(const b200fhe_shard* s, int* rank, int* world, int* device, int* nccl_version) {
    if (!s) return b200_set_err(B200FHE_EINVAL, "shard_info: null");
    if (rank) *rank = s->rank;
    if (world) *world = s->world;
    if (device) *device = s->device;
    if (nccl_version) { *nccl_version = 0; NcclApi* api = s->world > 1 ? nccl_api() : nullptr; if (api) api->GetVersion(nccl_version); }
    return 0;
}

// ---- the sharded ModUp plan ----------------------------------------------------------------------------------------------
extern "C" int b200fhe_modup_sharded_destroy(b200fhe_modup_plan* p) {
    if (!p) return 0;
    if (p->sh) cudaSetDevice(p->sh->device);
    for (b200fhe_bconv* c : p->conv) b200fhe_bconv_destroy(c);
    if (p->d_mods_scaled) cudaFree(p->d_mods_scaled);
    if (p->d_x) cudaFree(p->d_x);
    if (p->d_yfull) cudaFree(p->d_yfull);
    if (p->d_up_map) cudaFree(p->d_up_map);
    if (p->ev_scaled) cudaEventDestroy(p->ev_scaled);
    for (cudaEvent_t e : p->ev_gathered) cudaEventDestroy(e);
    if (p->ctx) b200fhe_ctx_destroy(p->ctx);
    delete p;
    return 0;
}

// moduli: the n_q ciphertext moduli followed by the n_p special primes, GLOBAL order (the same array on every rank).
extern "C" int b200fhe_modup_sharded_create(b200fhe_shard* sh, int log_n, const uint64_t* moduli, int n_q, int n_p, int dnum,
                                            b200fhe_modup_plan** out) {
    if (!sh || !moduli || !out || n_q < 1 || n_p < 1 || dnum < 1 || n_q % dnum) return b200_set_err(B200FHE_EINVAL, "modup_sharded_create: bad arguments");
    const int total = n_q + n_p, alpha = n_q / dnum, world = sh->world, rank = sh->rank;
    if (alpha > 64) return b200_set_err(B200FHE_EUNSUPPORTED, "modup_sharded_create: digits of more than 64 limbs");
    b200fhe_modup_plan* p = new b200fhe_modup_plan();
    p->sh = sh; p->n_q = n_q; p->n_p = n_p; p->dnum = dnum; p->alpha = alpha; p->total = total;
    std::vector<int> order, pos(total);
    for (int g = 0; g < total; g++)
        if (g % world == rank) { order.push_back(g); p->own_all.push_back(g); if (g < n_q) p->own_q.push_back(g); }
    for (int g = 0; g < total; g++)
        if (g % world != rank) order.push_back(g);
    std::vector<uint64_t> perm(total);
    for (int i = 0; i < total; i++) { perm[i] = moduli[order[i]]; pos[order[i]] = i; }
    int rc = b200fhe_ctx_create(log_n, perm.data(), total, sh->device, &p->ctx);
    if (rc) { b200fhe_modup_sharded_destroy(p); return rc; }
    b200fhe_ctx* c = p->ctx;
    const size_t N = c->N;
    p->counts.assign(dnum, std::vector<int>(world, 0));
    p->digit_row0.assign(dnum, 0); p->digit_rows.assign(dnum, 0); p->below.assign(dnum, 0);
    std::vector<B200Mod> ms(c->h_mods.begin(), c->h_mods.begin() + total);
    for (int d = 0; d < dnum && !rc; d++) {
        std::vector<int> src, dst;                                      // context positions
        for (int r = 0; r < world; r++)
            for (int g = d * alpha; g < (d + 1) * alpha; g++)
                if (g % world == r) { src.push_back(pos[g]); p->counts[d][r]++; }
        int row0 = -1, rows = 0, below = 0;
        for (size_t i = 0; i < p->own_all.size(); i++) {
            const int g = p->own_all[i];
            if (g >= d * alpha && g < (d + 1) * alpha) { if (row0 < 0) row0 = (int)i; rows++; }
            else { dst.push_back(pos[g]); if (g < d * alpha) below++; }
        }
        p->digit_row0[d] = row0 < 0 ? 0 : row0; p->digit_rows[d] = rows; p->below[d] = below;
        b200fhe_bconv* bc = nullptr;
        if (!dst.empty()) rc = b200fhe_bconv_create(c, src.data(), alpha, dst.data(), (int)dst.size(), &bc);
        p->conv.push_back(bc);
        // scale step of the digit folded into the inverse transform of its owned limbs (as keyswitch.cu does on one GPU)
        for (int g = d * alpha; g < (d + 1) * alpha; g++) {
            if (g % world != rank) continue;
            const uint64_t q = moduli[g];
            uint64_t h = 1;
            for (int j = d * alpha; j < (d + 1) * alpha; j++)
                if (j != g) h = nt::mulmod(h, moduli[j] % q, q);
            const uint64_t iv = nt::invmod(h, q);
            B200Mod& m = ms[pos[g]];
            m.ninv = nt::mulmod(m.ninv, iv, q);  m.ninv_s = nt::shoup(m.ninv, q);
            m.iw1n = nt::mulmod(m.iw1n, iv, q);  m.iw1n_s = nt::shoup(m.iw1n, q);
        }
    }
    if (rc) { b200fhe_modup_sharded_destroy(p); return rc; }
    cudaError_t e = cudaSetDevice(sh->device);
    if (e == cudaSuccess) e = cudaMalloc(&p->d_mods_scaled, sizeof(B200Mod) * total);
    if (e == cudaSuccess) e = cudaMemcpy(p->d_mods_scaled, ms.data(), sizeof(B200Mod) * total, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMalloc(&p->d_x, sizeof(uint64_t) * (p->own_q.empty() ? 1 : p->own_q.size()) * N);
    if (e == cudaSuccess) e = cudaMalloc(&p->d_yfull, sizeof(uint64_t) * (size_t)n_q * N);
    {
        const size_t n_all = p->own_all.size();
        for (int d = 0; d < dnum; d++)
            for (size_t r = 0; r < n_all; r++)
                if (p->conv[d] && (r < (size_t)p->below[d] || r >= (size_t)p->below[d] + (size_t)p->digit_rows[d]))
                    p->up_map.push_back(make_uint2((unsigned)(d * n_all + r), (unsigned)r));
        if (!p->up_map.empty()) {
            if (e == cudaSuccess) e = cudaMalloc(&p->d_up_map, sizeof(uint2) * p->up_map.size());
            if (e == cudaSuccess) e = cudaMemcpy(p->d_up_map, p->up_map.data(), sizeof(uint2) * p->up_map.size(), cudaMemcpyHostToDevice);
        }
    }
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_scaled, cudaEventDisableTiming);
    p->ev_gathered.assign(dnum, nullptr);
    for (int d = 0; d < dnum && e == cudaSuccess; d++) e = cudaEventCreateWithFlags(&p->ev_gathered[d], cudaEventDisableTiming);
    if (e != cudaSuccess) {
        b200fhe_modup_sharded_destroy(p);
        return b200_set_err(e == cudaErrorMemoryAllocation ? B200FHE_ENOMEM : B200FHE_ECUDA, "modup_sharded_create: %s", cudaGetErrorString(e));
    }
    *out = p;
    return 0;
}

extern "C" int b200fhe_modup_sharded_info(const b200fhe_modup_plan* p, int* n_own_q, int* n_own_total, int* own_limbs) {
    if (!p) return b200_set_err(B200FHE_EINVAL, "modup_sharded_info: null plan");
    if (n_own_q) *n_own_q = (int)p->own_q.size();
    if (n_own_total) *n_own_total = (int)p->own_all.size();
    if (own_limbs) for (size_t i = 0; i < p->own_all.size(); i++) own_limbs[i] = p->own_all[i];
    return 0;
}

// step 1: d_y_own [n_own_q][N] = INTT(d_c_own) (D_d / q_j)^-1 mod q_j, the rank's rows of every digit's scaled source matrix
extern "C" int b200fhe_modup_sharded_scale(b200fhe_modup_plan* p, const uint64_t* d_c_own, uint64_t* d_y_own, void* stream) {
    if (!p || (!p->own_q.empty() && (!d_c_own || !d_y_own))) return b200_set_err(B200FHE_EINVAL, "modup_sharded_scale: null argument");
    if (p->own_q.empty()) return 0;
    CK(cudaSetDevice(p->sh->device));
    const size_t N = p->ctx->N, n = p->own_q.size();
    if (d_y_own != d_c_own) {
        // out of place where the two-pass kernels serve the size (no device copy of the owned limbs first)
        const int rc = b200_nwt_backward_mods_oop(p->ctx, d_c_own, d_y_own, 1, n, 0, p->d_mods_scaled, stream);
        if (rc <= 0) return rc;
        CK(cudaMemcpyAsync(d_y_own, d_c_own, sizeof(uint64_t) * n * N, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    }
    return b200_nwt_backward_mods(p->ctx, d_y_own, 1, n, 0, p->d_mods_scaled, stream);
}

// step 3 of one digit: d_y_full [alpha][N] = the digit's scaled limbs in gather (rank-major) order
static bool merged_transform(const b200fhe_modup_plan* p) {
    static const bool no_merge = getenv("B200FHE_KS_NO_MERGE") != nullptr;
    return !no_merge && p->d_up_map && p->ctx->log_n >= 12 && p->up_map.size() <= 65535;
}
static int transform_all(b200fhe_modup_plan* p, uint64_t* d_up_own, void* stream) {
    if (!merged_transform(p)) return 0;
    const int rc = b200_nwt_forward_map(p->ctx, d_up_own, p->d_up_map, p->up_map.data(), p->up_map.size(), stream);
    return rc > 0 ? b200_set_err(B200FHE_EUNSUPPORTED, "modup_sharded: scatter transform unavailable") : rc;
}
// with_ntt = false: contraction only (the merged transform follows the last digit)
static int raise_digit(b200fhe_modup_plan* p, int d, const uint64_t* d_c_own, const uint64_t* d_y_full, uint64_t* d_up_own, void* stream,
                       bool with_ntt = true) {
    b200fhe_ctx* c = p->ctx;
    const size_t N = c->N, n_all = p->own_all.size();
    uint64_t* upd = d_up_own + (size_t)d * n_all * N;
    const size_t below = (size_t)p->below[d], rows = (size_t)p->digit_rows[d], above = n_all - below - rows;
    if (rows)   // the digit's own limbs pass through (rows [below, below + rows) of the output are rows digit_row0.. of the input)
        CK(cudaMemcpyAsync(upd + below * N, d_c_own + (size_t)p->digit_row0[d] * N, sizeof(uint64_t) * rows * N, cudaMemcpyDeviceToDevice,
                           (cudaStream_t)stream));
    if (!p->conv[d]) return 0;
    if (!with_ntt) {
        const int one = b200_bconv_contract_skip(p->conv[d], d_y_full, upd, below, rows, stream);
        if (one <= 0) return one;
        if (below) RC(b200fhe_bconv_contract(p->conv[d], d_y_full, upd, 0, below, stream));
        if (above) RC(b200fhe_bconv_contract(p->conv[d], d_y_full, upd + (below + rows) * N, below, above, stream));
        return 0;
    }
    if (below) {
        RC(b200fhe_bconv_contract(p->conv[d], d_y_full, upd, 0, below, stream));
        RC(b200fhe_nwt_forward(c, upd, 1, below, 0, stream));
    }
    if (above) {
        RC(b200fhe_bconv_contract(p->conv[d], d_y_full, upd + (below + rows) * N, below, above, stream));
        RC(b200fhe_nwt_forward(c, upd + (below + rows) * N, 1, above, below + rows, stream));
    }
    return 0;
}

// steps 3 of every digit given the gathered digits (a caller with its own transport, and the single-GPU emulation of the tests)
extern "C" int b200fhe_modup_sharded_raise(b200fhe_modup_plan* p, const uint64_t* d_c_own, const uint64_t* const* d_y_full, uint64_t* d_up_own,
                                           void* stream) {
    if (!p || !d_y_full || !d_up_own || (!p->own_q.empty() && !d_c_own)) return b200_set_err(B200FHE_EINVAL, "modup_sharded_raise: null argument");
    CK(cudaSetDevice(p->sh->device));
    const bool merged = merged_transform(p);
    for (int d = 0; d < p->dnum; d++) RC(raise_digit(p, d, d_c_own, d_y_full[d], d_up_own, stream, !merged));
    return transform_all(p, d_up_own, stream);
}

// The whole sharded ModUp of one polynomial.  d_c_own [n_own_q][N] (NTT domain, the rank's limbs of Q ascending) ->
// d_up_own [dnum][n_own_total][N] (NTT domain, the rank's limbs of Q u P ascending, for every digit).
extern "C" int b200fhe_modup_sharded(b200fhe_modup_plan* p, const uint64_t* d_c_own, uint64_t* d_up_own, void* stream) {
    if (!p || !d_up_own || (!p->own_q.empty() && !d_c_own)) return b200_set_err(B200FHE_EINVAL, "modup_sharded: null argument");
    b200fhe_shard* sh = p->sh;
    CK(cudaSetDevice(sh->device));
    cudaStream_t s = (cudaStream_t)stream;
    const size_t N = p->ctx->N, a = (size_t)p->alpha;
    RC(b200fhe_modup_sharded_scale(p, d_c_own, p->d_x, stream));
    const bool merged = merged_transform(p);
    if (sh->world == 1) {
        for (int d = 0; d < p->dnum; d++) RC(raise_digit(p, d, d_c_own, p->d_x + (size_t)d * a * N, d_up_own, stream, !merged));
        return transform_all(p, d_up_own, stream);
    }
    NcclApi* api = nccl_api();
    if (!api || !sh->comm) return b200_set_err(B200FHE_EUNSUPPORTED, "modup_sharded: no NCCL communicator on this shard (use the scale / raise halves)");
    CK(cudaEventRecord(p->ev_scaled, s));
    CK(cudaStreamWaitEvent(sh->comm_stream, p->ev_scaled, 0));
    for (int d = 0; d < p->dnum; d++) {
        uint64_t* yf = p->d_yfull + (size_t)d * a * N;
        const uint64_t* send = p->d_x + (size_t)p->digit_row0[d] * N;
        const std::vector<int>& cnt = p->counts[d];
        bool equal = true;
        for (int r = 1; r < sh->world; r++) equal = equal && cnt[r] == cnt[0];
        if (equal) {
            NC(api, api->AllGather(send, yf, (size_t)cnt[0] * N, kNcclUint64, sh->comm, sh->comm_stream));
        } else {
            NC(api, api->GroupStart());
            size_t off = 0;
            for (int r = 0; r < sh->world; r++) {
                if (cnt[r]) NC(api, api->Broadcast(send, yf + off * N, (size_t)cnt[r] * N, kNcclUint64, r, sh->comm, sh->comm_stream));
                off += (size_t)cnt[r];
            }
            NC(api, api->GroupEnd());
        }
        CK(cudaEventRecord(p->ev_gathered[d], sh->comm_stream));
    }
    for (int d = 0; d < p->dnum; d++) {
        CK(cudaStreamWaitEvent(s, p->ev_gathered[d], 0));
        RC(raise_digit(p, d, d_c_own, p->d_yfull + (size_t)d * a * N, d_up_own, stream, !merged));
    }
    return transform_all(p, d_up_own, stream);
}
