// Host-side context shared by the CUDA translation units and the C ABI.
// Mirrors what the reference harness assembles by hand from Phantom types
// (reliability_test/ntt_test.cu:44-69: CoeffModulus::Create -> DModulus[] ->
// DNTTTable.init/set per limb): moduli, Barrett/Shoup constants and the
// bit-reversed twiddle tables, resident on one device.
#pragma once
#include <atomic>
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>
#include <cuda_runtime.h>
#include "modarith.cuh"
#include "ntt_phases.cuh"

namespace b200 {

// rows of 256 coefficients per row-pass CTA (one row per half-warp); also the tile size of the row table
#ifndef B200_ROWS_PER_CTA
#define B200_ROWS_PER_CTA 8
#endif
constexpr int kRowsPerCta = B200_ROWS_PER_CTA;
constexpr int kRowTileEntries = kRowsPerCta * 256;   // ulonglong2 entries per tile: A[rows][16] + B[15][rows*16]

struct PassArgs {
    u64* data;                // [n_limbs][N]
    const B200Mod* mods;      // [ctx.L]
    const ulonglong2* tw;     // [ctx.L][N] interleaved (w, w_shoup), forward or inverse table (SEAL layout)
    const ulonglong2* rtw;    // [ctx.L][N] the same table's entries [N/256, N) re-tiled for the row pass
    u32 N;
    u32 L;                    // limbs per polynomial in this call
    u32 start_mod;
    u32 n_poly = 1;           // polynomials in this launch (grid.x = tiles * n_poly)
    u32 poly0 = 0;            // row passes: first polynomial of this launch (an odd last polynomial of a batch runs the one-poly form)
    // optional scatter list (n_poly = 1): entry y of the launch transforms data limb map[y].x with modulus map[y].y -- limbs
    // that are neither contiguous nor in modulus order (the converted limbs of every key-switch digit) in ONE launch
    const uint2* map = nullptr;
    // optional source of the inverse transform's FIRST pass (same layout as data): the rows are read from src and written to data,
    // so a caller that must keep its input (the key-switch keeps c) needs no device copy first.  Plain transform, one polynomial per CTA.
    const u64* src = nullptr;
    // protected (ABFT) forward transform only
    const u64* omega = nullptr;     // [ctx.L][N] output weights, 30-bit halves packed (lo | hi << 32)
    u64* abft = nullptr;            // [n_limbs of the call][2][4] digit-sum counters
    const InjEntry* inj = nullptr;  // device copy of the injection plan
    u32 n_inj = 0;
    u32 limb0 = 0;                  // call-level index of this launch's first limb
};

// mask: 1 = first pass only, 2 = second pass only, 3 = whole transform
// mode: 0 = any modulus, 1 = q >= 2^36 (BIG), 2 = q = 2^bits - delta (NEAR)
int launch_ntt(bool inverse, int big, const PassArgs& a, int logn, u32 n_limbs, int mask, cudaStream_t s);
// protected forward transform (a.omega / a.abft set); counters must be zero on entry
int launch_ntt_protected(int big, const PassArgs& a, int logn, u32 n_limbs, cudaStream_t s, int mask = 3);
// protected inverse transform: a.tw / a.rtw are the inverse tables; side 0 = evaluation-side sum, side 1 = coefficient-side sum
int launch_ntt_protected_inverse(int big, const PassArgs& a, int logn, u32 n_limbs, cudaStream_t s);
// verdict[i] = (input checksum != output checksum); sums (optional) [n][2]; zeroes the counters
void launch_abft_finalize(u64* counters, const B200Mod* mods, u32 L, u32 start_mod, u32 n_limbs, u32* verdict, u64* sums, cudaStream_t s);

// ct[k] -> [P][L][N], pt[k] -> [L][N] (host arrays of T device pointers), out [P][L][N]
int launch_pmult_moddown(const u64* const* ct, const u64* const* pt, size_t T, const u64* conv, u64* out, const B200Mod* mods,
                         const ulonglong2* pinv, u32 N, u32 L, cudaStream_t s, const u64* c_own = nullptr, u32 alpha = 0);
int launch_pmult(const u64* const* ct, const u64* const* pt, size_t T, u64* out, const B200Mod* mods, u32 N, u32 P, u32 L,
                 u32 start_mod, int accumulate, cudaStream_t s, unsigned long long* launches);

int launch_bconv_scale(const u64* x, u64* y, const B200Mod* mods, const u32* src_idx, const ulonglong2* inv, u32 N, u32 j0, u32 count,
                       cudaStream_t s);
int launch_bconv_contract(const u64* y, u64* out, const B200Mod* mods, const u32* dst_idx, const u64* hat, u32 N, u32 m, u32 K, u32 k0,
                          u32 kcount, cudaStream_t s);
// exact-CRT mode: takes the e P overshoot of the fast conversion out of out [K][N] (hatw [m][m] words of P/p_j, ePw [m][m] words of e P, ePq [K][m])
int launch_bconv_exact_fix(const u64* y, u64* out, u32 N, u32 m, u32 K, const u64* hatw, const u64* ePw, const u64* ePq, const B200Mod* mods,
                           const u32* dst_idx, cudaStream_t s);
// tensor-core (byte-split, mma.sync u8) contraction for source bases of at most 16 moduli; btab [8][ceil(m/4)][K][8] fragments
int launch_bconv_contract_tc(const u64* y, u64* out, const B200Mod* mods, const u32* dst_idx, const uint2* btab, u32 N, u32 m, u32 K,
                             u32 k0, u32 kcount, cudaStream_t s);
// tcgen05 / TMEM contraction (bconv_tc5.cu); returns -1 when the shape is not served there (caller falls back to launch_bconv_contract_tc)
int launch_bconv_contract_tc5(const u64* y, u64* out, const B200Mod* mods, const u32* dst_idx, const uint4* btab5, u32 N, u32 m, u32 K,
                              u32 k0, u32 kcount, cudaStream_t s);
int launch_bconv_contract_tc5_skip(const u64* y, u64* out, const B200Mod* mods, const u32* dst_idx, const uint4* btab5, u32 N, u32 m, u32 K,
                                   u32 skip_at, u32 skip_rows, cudaStream_t s);
struct BconvInj {          // device copy of b200fhe_bconv_inject
    u32 stage, limb;
    u64 coeff, mask;
};
constexpr int kBconvMaxInj = 8;
struct BconvInjList {
    BconvInj e[kBconvMaxInj];
    u32 n;
};
// checksum-carrying flavour of the tcgen05 contraction: cin [m][2] / cout [K][2] sums of the 32-bit halves of the words read / stored
int launch_bconv_contract_tc5_prot(const u64* y, u64* out, const B200Mod* mods, const u32* dst_idx, const uint4* btab5, u32 N, u32 m, u32 K,
                                   u32 k0, u32 kcount, cudaStream_t s, u64* cin, u64* cout, const BconvInjList& inj);
// checksum-carrying ModUp of one digit (scale + tensor-core contraction + finalize); counters: 6 m + 2 K words of scratch
int launch_bconv_modup_protected(const u64* x, u64* y, u64* out, const B200Mod* mods, const u32* src_idx, const u32* dst_idx,
                                 const ulonglong2* inv, const u64* hat, const uint2* btab, u32 N, u32 m, u32 K, u64* counters,
                                 u64* next_counters, u32* verdict, u64* sums, const void* inj_host, u32 n_inj, cudaStream_t s, unsigned long long* launches,
                                 u64* out_hi = nullptr, u32 k_split = 0, const uint4* btab5 = nullptr);

// Opt-in dynamic shared memory above 48 KiB, set once per kernel instantiation (the kernel is a non-type template argument: one
// static per instantiation, not per function-pointer type) and per device (the attribute belongs to the device's context).
template <auto Kernel>
static inline cudaError_t set_smem_once(size_t smem) {
    static std::atomic<unsigned long long> done{0};
    if (smem <= 48 * 1024) return cudaSuccess;
    int dev = 0;
    cudaGetDevice(&dev);
    const unsigned long long bit = 1ull << (dev & 63);
    if (done.load(std::memory_order_relaxed) & bit) return cudaSuccess;
    const cudaError_t rc = cudaFuncSetAttribute(Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (rc == cudaSuccess) done.fetch_or(bit, std::memory_order_relaxed);
    return rc;
}

}  // namespace b200

constexpr unsigned kAbftSlots = 4;   // protected calls that may be in flight per context (each has its own counters)

struct b200fhe_ctx {
    int log_n = 0;
    size_t N = 0;
    int L = 0;                       // number of moduli held
    int device = 0;
    std::vector<uint64_t> moduli;
    std::vector<uint64_t> psi;       // minimal primitive 2N-th roots
    std::vector<B200Mod> h_mods;
    std::vector<uint8_t> has_inverse;   // per modulus: inverse tables present
    B200Mod* d_mods = nullptr;       // [L]
    ulonglong2* d_tw = nullptr;      // [L][N] forward (w, w')
    ulonglong2* d_itw = nullptr;     // [L][N] inverse
    ulonglong2* d_rtw = nullptr;     // [L][N] forward, row-pass tiles (N >= 2^12)
    ulonglong2* d_irtw = nullptr;    // [L][N] inverse, row-pass tiles
    // ABFT (built lazily by the first protected call)
    uint64_t* d_omega = nullptr;     // [L][N]
    std::vector<uint8_t> has_omega;  // per modulus
    uint64_t* d_abft = nullptr;      // counters: ring of kAbftSlots slices of 16 abft_limbs words, grown on demand
    size_t abft_limbs = 0;
    unsigned abft_next = 0;          // next ring slot (one per protected call)
    unsigned abft_dirty = 0;         // bit s: slot s may hold non-zero counters (a call failed between launch and finalize)
    InjEntry* d_inj = nullptr;       // injection plan staging, kAbftSlots x B200FHE_MAX_INJECT entries
    // host-buffer entry points: pinned staging + device scratch, grown on demand
    uint64_t* h_stage = nullptr;
    uint64_t* d_stage = nullptr;
    size_t stage_words = 0, hstage_words = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
    std::vector<cudaEvent_t> ev;
};

// Base-conversion plan: constants of one (source base -> target moduli) pair, resident on the context's device.
struct b200fhe_bconv {
    b200fhe_ctx* ctx = nullptr;
    int m = 0, K = 0;
    u32* d_src = nullptr;            // [m] modulus indices
    u32* d_dst = nullptr;            // [K]
    ulonglong2* d_inv = nullptr;     // [m] (P/p_j)^-1 mod p_j with Shoup quotient
    u64* d_hat = nullptr;            // [K][m] (P/p_j) mod q_k, 30-bit halves packed
    uint2* d_btab = nullptr;         // m <= 16: the same constants as byte-split tensor-core B fragments (bconv_kernels.cu)
    uint4* d_btab5 = nullptr;        // m <= 16: the same constants as tcgen05 K-major B rows [K][8][16] x 16 bytes (bconv_tc5.cu)
    int mode = 0;                    // 0: fast (HPS) conversion, 1: exact CRT (motivation/baseConv.py:67-83)
    u64* d_hatw = nullptr;           // exact mode: [m][m] words of P/p_j
    u64* d_ePw = nullptr;            //             [m][m] words of e P, e = 0..m-1
    u64* d_ePq = nullptr;            //             [K][m]  e P mod q_k
    u64* d_y = nullptr;              // [m][N] scratch for the single-GPU path
    u64* d_cnt = nullptr;            // checksum counters of the protected path: two sets of 6 m + 2 K words, used alternately
    int cnt_set = 0;
    u64* d_io = nullptr;             // host-buffer path staging, grown on demand
    size_t io_words = 0;
};
