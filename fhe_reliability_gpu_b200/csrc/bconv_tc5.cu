// Base-conversion contraction on the 5th-generation tensor cores: tcgen05.mma.kind::i8 with accumulators in TMEM.
//
// Same byte-split formulation as k_bconv_contract_tc (bconv_kernels.cu; semantics rfhe_framewk/src/baseConv.py:10-40): with
// k-index (j, a) = byte a of y_j and the constants Toeplitz-expanded, B[(t, d)][(j, a)] = byte (d - a) of hat_jt, a u8 x u8 -> s32
// product gives per output the 16 digit sums S_d (< 2^23) and out = (sum_d S_d 2^(8d)) mod q_t.  What changes is WHO multiplies:
// one thread issues tcgen05.mma (M = 128 coefficients, N = 256 = 16 targets x 16 digits, K = 32 bytes per instruction), the
// accumulators land in tensor memory, and the 16 warps of the CTA do nothing but the epilogue (tcgen05.ld of a thread's 16 digit
// sums -> recombination -> one 128-bit Barrett -> coalesced store) of tile t - 1 while the tensor core works on tile t: the two
// phases that mma.sync serialised inside every warp (4.6 k cycles per group measured against ~2 k for the longer phase alone,
// profiles/r1j_bconv_tc_summary.txt) now overlap.
//
// Shared memory (canonical K-major, no swizzle; in 16-byte units ((8, n), 2) : ((1, SBO), LBO)):
//   A tile  [128 rows][8 m bytes]   byte (r, k) at (k / 16) * 2048 + r * 16 + k % 16   SBO 128, LBO 2048   (two buffers)
//   B chunk [256 cols][128 bytes]   byte (n, k) at (k / 16) * 4096 + n * 16 + k % 16   SBO 128, LBO 4096   (up to 3 chunks resident)
// The y words need no repacking beyond this placement: the 8 little-endian bytes of y[j][i] ARE bytes 8 j .. 8 j + 7 of row i.
// TMEM: all 512 columns = two accumulator buffers of 256 columns (one CTA per SM, persistent over the coefficient tiles).
// Layout and descriptor encodings validated in isolation by profiles/microbench/tc5/umma_i8_test.cu.
#include "ctx.hpp"
#include <cstdlib>

namespace b200 {

constexpr int kT5Rows = 128, kT5Chunk = 16, kT5Threads = 512, kT5MaxChunks = 3;
constexpr u32 kT5ABytes = kT5Rows * 128, kT5BBytes = 256 * 128;

__device__ __forceinline__ u32 t5_smem(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ u64 t5_desc(u32 addr, u32 lbo, u32 sbo) {
    return (u64)((addr >> 4) & 0x3fff) | ((u64)((lbo >> 4) & 0x3fff) << 16) | ((u64)((sbo >> 4) & 0x3fff) << 32) | ((u64)1 << 46);
}
// c_format S32 (2) bits [4,6), a / b format UINT8 (0), both K-major, n_dim = 256 >> 3 bits [17,23), m_dim = 128 >> 4 bits [24,29)
constexpr u32 kT5Idesc = (2u << 4) | (32u << 17) | (8u << 24);

__device__ __forceinline__ void t5_digits_value(const u32 (&S)[16], u64& lo, u64& hi) {     // sum_d S_d 2^(8d), S_d < 2^23
    u64 W[4];
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const u32 x = S[4 * r] + (S[4 * r + 1] << 8), yv = S[4 * r + 2] + (S[4 * r + 3] << 8);
        W[r] = (u64)x + ((u64)yv << 16);
    }
    lo = W[0] + (W[1] << 32);
    hi = (W[1] >> 32) + W[2] + (W[3] << 32) + (lo < W[0]);
}

__device__ __forceinline__ void t5_red_add64(u64* p, u64 v) { asm volatile("red.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }

struct T5Smem {
    unsigned char a[2][kT5ABytes];
    unsigned char b[kT5MaxChunks][kT5BBytes];
    ulonglong4 q[kT5MaxChunks * kT5Chunk];
    u64 bar[2];
    u32 tmem;
    u64 so[kT5Threads / 32][kT5MaxChunks][8];      // PROT: per-warp output sums, [4 targets of the warp][low / high half]
};

// PROT (checksum-carrying flavour): cin [m][2] += sums of the low / high 32-bit halves of the y words as this
// kernel loads them (nullptr: not this launch), cout [K][2] += the same sums of the outputs as stored; stage-2 faults of `inj` are
// XORed into the stored (and summed) outputs.  The sums ride on data already in registers: no extra pass over memory.
template <bool PROT>
__global__ void __launch_bounds__(kT5Threads, 1)
k_bconv_contract_tc5(const u64* __restrict__ y, u64* __restrict__ out, const B200Mod* mods, const u32* dst_idx,
                     const uint4* __restrict__ btab5 /*[K][8][16] 16-byte entries*/, u32 N, u32 m, u32 k0, u32 kcount, u64* cin, u64* cout,
                     const __grid_constant__ BconvInjList inj, u32 skip_at, u32 skip_rows) {
    extern __shared__ __align__(1024) unsigned char smraw[];
    T5Smem& s = *reinterpret_cast<T5Smem*>(smraw);
    const u32 tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const u32 nch = (kcount + kT5Chunk - 1) / kT5Chunk, ks = (m + 3) / 4, n_tiles = N / kT5Rows;
    // constants: B chunks (zero columns for targets past kcount), per-target (q, ratio)
    for (u32 idx = tid; idx < nch * (kT5BBytes / 16); idx += kT5Threads) {
        const u32 c = idx / (kT5BBytes / 16), r = idx % (kT5BBytes / 16), kc = r >> 8, tl = (r >> 4) & 15, d = r & 15;
        const u32 t = c * kT5Chunk + tl;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (t < kcount) v = __ldg(btab5 + ((size_t)(k0 + t) * 8 + kc) * 16 + d);
        *reinterpret_cast<uint4*>(s.b[c] + kc * 4096 + tl * 256 + d * 16) = v;
    }
    for (u32 t = tid; t < nch * kT5Chunk; t += kT5Threads) {
        const B200Mod& mk = mods[dst_idx[k0 + (t < kcount ? t : kcount - 1)]];
        s.q[t] = make_ulonglong4(mk.q, mk.ratio0, mk.ratio1, 0);
    }
    if (PROT)
        for (u32 k = tid; k < (kT5Threads / 32) * kT5MaxChunks * 8; k += kT5Threads) (&s.so[0][0][0])[k] = 0;
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(t5_smem(&s.bar[0])));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(t5_smem(&s.bar[1])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(t5_smem(&s.tmem)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const u32 tm = s.tmem;
    const u32 quarter = warp & 3, cg = warp >> 2;               // TMEM lanes 32 quarter .. + 31; targets 4 cg .. 4 cg + 3 of a chunk

    // epilogue of one (tile, chunk) item whose accumulators sit in TMEM buffer `buf`
    auto epilogue = [&](u32 item, u32 mt, u32 c) {
        const u32 buf = item & 1, parity = (item >> 1) & 1;
        asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(
                         t5_smem(&s.bar[buf])),
                     "r"(parity)
                     : "memory");
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const u32 row = mt * kT5Rows + quarter * 32 + lane;
        // one load brings the thread's 64 columns = the 16 digit sums of its four targets; the four recombination + Barrett
        // chains below are independent, so the scheduler interleaves them (targets past kcount are zero columns: computed, not stored)
        u32 S[4][16];
        const u32 ta = tm + ((quarter * 32u) << 16) + buf * 256 + cg * 64;
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
            "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,"
            "%32,%33,%34,%35,%36,%37,%38,%39,%40,%41,%42,%43,%44,%45,%46,%47,%48,%49,%50,%51,%52,%53,%54,%55,%56,%57,%58,%59,%60,%61,%62,%63}, [%64];"
            : "=r"(S[0][0]), "=r"(S[0][1]), "=r"(S[0][2]), "=r"(S[0][3]), "=r"(S[0][4]), "=r"(S[0][5]), "=r"(S[0][6]), "=r"(S[0][7]),
              "=r"(S[0][8]), "=r"(S[0][9]), "=r"(S[0][10]), "=r"(S[0][11]), "=r"(S[0][12]), "=r"(S[0][13]), "=r"(S[0][14]), "=r"(S[0][15]),
              "=r"(S[1][0]), "=r"(S[1][1]), "=r"(S[1][2]), "=r"(S[1][3]), "=r"(S[1][4]), "=r"(S[1][5]), "=r"(S[1][6]), "=r"(S[1][7]),
              "=r"(S[1][8]), "=r"(S[1][9]), "=r"(S[1][10]), "=r"(S[1][11]), "=r"(S[1][12]), "=r"(S[1][13]), "=r"(S[1][14]), "=r"(S[1][15]),
              "=r"(S[2][0]), "=r"(S[2][1]), "=r"(S[2][2]), "=r"(S[2][3]), "=r"(S[2][4]), "=r"(S[2][5]), "=r"(S[2][6]), "=r"(S[2][7]),
              "=r"(S[2][8]), "=r"(S[2][9]), "=r"(S[2][10]), "=r"(S[2][11]), "=r"(S[2][12]), "=r"(S[2][13]), "=r"(S[2][14]), "=r"(S[2][15]),
              "=r"(S[3][0]), "=r"(S[3][1]), "=r"(S[3][2]), "=r"(S[3][3]), "=r"(S[3][4]), "=r"(S[3][5]), "=r"(S[3][6]), "=r"(S[3][7]),
              "=r"(S[3][8]), "=r"(S[3][9]), "=r"(S[3][10]), "=r"(S[3][11]), "=r"(S[3][12]), "=r"(S[3][13]), "=r"(S[3][14]), "=r"(S[3][15])
            : "r"(ta));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        u64 res[4];
#pragma unroll
        for (u32 tl = 0; tl < 4; tl++) {
            const u32 t = c * kT5Chunk + cg * 4 + tl;
            u64 lo, hi;
            t5_digits_value(S[tl], lo, hi);
            const ulonglong4 mk = s.q[t];
            res[tl] = barrett128_raw(lo, hi, mk.x, mk.y, mk.z);
        }
        if (PROT) {
            for (u32 e = 0; e < inj.n; e++)                       // a faulty reduction: the corrupted output is stored AND summed
                if (inj.e[e].stage == 2 && inj.e[e].coeff == row) {
#pragma unroll
                    for (u32 tl = 0; tl < 4; tl++)
                        if (inj.e[e].limb == k0 + c * kT5Chunk + cg * 4 + tl) res[tl] ^= inj.e[e].mask;
                }
            // sums over the warp's 32 rows of the 8 values (4 targets x 2 halves), transposing while reducing: each round halves
            // the values a lane carries, 4 + 2 + 1 + 1 + 1 = 9 64-bit shuffles instead of 8 x 5
            u64 v[8];
#pragma unroll
            for (u32 tl = 0; tl < 4; tl++) {
                const bool live = c * kT5Chunk + cg * 4 + tl < kcount;
                v[2 * tl] = live ? (res[tl] & 0xffffffffull) : 0;
                v[2 * tl + 1] = live ? (res[tl] >> 32) : 0;
            }
            const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
            u64 w[4], x[2];
#pragma unroll
            for (int i = 0; i < 4; i++) w[i] = (b4 ? v[i + 4] : v[i]) + __shfl_xor_sync(0xffffffffu, b4 ? v[i] : v[i + 4], 16);
#pragma unroll
            for (int i = 0; i < 2; i++) x[i] = (b3 ? w[i + 2] : w[i]) + __shfl_xor_sync(0xffffffffu, b3 ? w[i] : w[i + 2], 8);
            u64 z = (b2 ? x[1] : x[0]) + __shfl_xor_sync(0xffffffffu, b2 ? x[0] : x[1], 4);
            z += __shfl_xor_sync(0xffffffffu, z, 2);
            z += __shfl_xor_sync(0xffffffffu, z, 1);
            // lane 4 e now holds value e = bit2 + 2 bit3 + 4 bit4 of the lane index: one lane per entry, plain read-add-write
            if ((lane & 3) == 0) s.so[warp][c][lane >> 2] += z;
        }
#pragma unroll
        for (u32 tl = 0; tl < 4; tl++) {
            const u32 t = c * kT5Chunk + cg * 4 + tl;
            // targets from skip_at on land skip_rows rows further down: the key-switch writes a digit's targets around the digit's
            // own limbs (which sit in the middle of the raised polynomial) in one launch
            if (t < kcount) out[(size_t)(t + (k0 + t >= skip_at ? skip_rows : 0u)) * N + row] = res[tl];
        }
    };

    // A tile: row r, limb j -> bytes 8 j .. 8 j + 7 of row r (rows fastest: coalesced 8-byte loads).  The NEXT tile's words are
    // loaded into registers before the epilogues of the current tile and stored into the other buffer after them, so the
    // global-memory latency hides under the epilogue work.
    u64 pre[4];
    u64 ci[4][2] = {};                                        // PROT: this thread's y sums; word k of a thread is always limb (tid >> 7) + 4 k
    auto load_tile = [&](u32 mt) {
#pragma unroll
        for (u32 k = 0; k < 4; k++) {
            const u32 idx = tid + k * kT5Threads;
            if (idx < kT5Rows * m) {
                pre[k] = __ldg(y + (size_t)(idx >> 7) * N + (size_t)mt * kT5Rows + (idx & (kT5Rows - 1)));
                if (PROT) { ci[k][0] += pre[k] & 0xffffffffull; ci[k][1] += pre[k] >> 32; }
            }
        }
    };
    auto store_tile = [&](unsigned char* A) {
#pragma unroll
        for (u32 k = 0; k < 4; k++) {
            const u32 idx = tid + k * kT5Threads;
            const u32 r = idx & (kT5Rows - 1), j = idx >> 7;
            if (idx < kT5Rows * m) *reinterpret_cast<u64*>(A + (j >> 1) * 2048 + r * 16 + (j & 1) * 8) = pre[k];
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");       // generic-proxy writes -> visible to the tensor core
    };
    u32 item = 0, prev_mt = 0, prev_c = 0, tcount = 0;
    if (blockIdx.x < n_tiles) { load_tile(blockIdx.x); store_tile(s.a[0]); }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    for (u32 mt = blockIdx.x; mt < n_tiles; mt += gridDim.x, tcount++) {
        unsigned char* A = s.a[tcount & 1];
        const u32 next = mt + gridDim.x;
        if (next < n_tiles) load_tile(next);
        for (u32 c = 0; c < nch; c++, item++) {
            if (tid == 0) {
                const u32 d_tmem = tm + (item & 1) * 256;
                for (u32 st = 0; st < ks; st++) {
                    const u64 ad = t5_desc(t5_smem(A) + st * 4096, 2048, 128);
                    const u64 bd = t5_desc(t5_smem(s.b[c]) + st * 8192, 4096, 128);
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
                                 "l"(ad), "l"(bd), "r"(kT5Idesc), "r"(st > 0 ? 1u : 0u)
                                 : "memory");
                }
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(t5_smem(&s.bar[item & 1])) : "memory");
            }
            if (item > 0) epilogue(item - 1, prev_mt, prev_c);
            prev_mt = mt; prev_c = c;
            // the other A buffer's last readers (the previous tile's MMAs) were all waited for by the epilogues above
            if (c == nch - 1 && next < n_tiles) store_tile(s.a[(tcount + 1) & 1]);
            // the TMEM buffer written two items from now is the one just drained: order the tcgen05.ld's before that MMA
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncthreads();
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
    }
    if (item > 0) epilogue(item - 1, prev_mt, prev_c);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (PROT) {
        // outputs: entry (c, cg, e) summed over the four row quarters (warps cg * 4 .. cg * 4 + 3), one global reduction per CTA and entry
        for (u32 k = tid; k < nch * 32; k += kT5Threads) {
            const u32 c = k >> 5, g4 = (k >> 3) & 3, e = k & 7, t = c * kT5Chunk + g4 * 4 + (e >> 1);
            const u64 v = s.so[g4 * 4][c][e] + s.so[g4 * 4 + 1][c][e] + s.so[g4 * 4 + 2][c][e] + s.so[g4 * 4 + 3][c][e];
            if (t < kcount) t5_red_add64(cout + (size_t)(k0 + t) * 2 + (e & 1), v);
        }
        if (cin) {
#pragma unroll
            for (u32 k = 0; k < 4; k++)
#pragma unroll
                for (u32 h = 0; h < 2; h++) {
                    u64 v = ci[k][h];
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                    const u32 j = (tid >> 7) + 4 * k;
                    if (lane == 0 && j < m) t5_red_add64(cin + (size_t)j * 2 + h, v);
                }
        }
    }
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
}

// Returns 0 when launched; -1 when this shape is not served here (the caller falls back to the mma.sync kernel).
template <bool PROT>
static int launch_t5(const u64* y, u64* out, const B200Mod* mods, const u32* dst_idx, const uint4* btab5, u32 N, u32 m, u32 k0, u32 kcount,
                     cudaStream_t st, u64* cin, u64* cout, const BconvInjList& inj, u32 skip_at = 0xffffffffu, u32 skip_rows = 0) {
    static const bool off = getenv("B200FHE_BCONV_NO_TC5") != nullptr;
    if (off || !btab5 || m > 16 || N % kT5Rows || N < 4096) return -1;
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    static int sm_count[64] = {};
    if (!sm_count[dev & 63]) cudaDeviceGetAttribute(&sm_count[dev & 63], cudaDevAttrMultiProcessorCount, dev);
    sms = sm_count[dev & 63];
    if (sms <= 0 || set_smem_once<k_bconv_contract_tc5<PROT>>(sizeof(T5Smem) + 1024) != cudaSuccess) return -1;
    const u32 n_tiles = N / kT5Rows;
    for (u32 off_k = 0; off_k < kcount; off_k += kT5MaxChunks * kT5Chunk) {
        const u32 kc = kcount - off_k < (u32)(kT5MaxChunks * kT5Chunk) ? kcount - off_k : (u32)(kT5MaxChunks * kT5Chunk);
        const u32 grid = n_tiles < (u32)sms ? n_tiles : (u32)sms;
        k_bconv_contract_tc5<PROT><<<grid, kT5Threads, sizeof(T5Smem) + 1024, st>>>(y, out + (size_t)off_k * N, mods, dst_idx, btab5, N, m, k0 + off_k,
                                                                                      kc, off_k ? nullptr : cin, cout, inj, skip_at, skip_rows);
    }
    return 0;
}

int launch_bconv_contract_tc5(const u64* y, u64* out, const B200Mod* mods, const u32* dst_idx, const uint4* btab5, u32 N, u32 m, u32 K,
                              u32 k0, u32 kcount, cudaStream_t st) {
    BconvInjList none;
    none.n = 0;
    return launch_t5<false>(y, out, mods, dst_idx, btab5, N, m, k0, kcount, st, nullptr, nullptr, none);
}

// every target of the plan in one pass, targets >= skip_at written skip_rows rows further down
int launch_bconv_contract_tc5_skip(const u64* y, u64* out, const B200Mod* mods, const u32* dst_idx, const uint4* btab5, u32 N, u32 m, u32 K,
                                   u32 skip_at, u32 skip_rows, cudaStream_t st) {
    BconvInjList none;
    none.n = 0;
    return launch_t5<false>(y, out, mods, dst_idx, btab5, N, m, 0, K, st, nullptr, nullptr, none, skip_at, skip_rows);
}

int launch_bconv_contract_tc5_prot(const u64* y, u64* out, const B200Mod* mods, const u32* dst_idx, const uint4* btab5, u32 N, u32 m, u32 K,
                                   u32 k0, u32 kcount, cudaStream_t st, u64* cin, u64* cout, const BconvInjList& inj) {
    if (!cout) return -1;
    return launch_t5<true>(y, out, mods, dst_idx, btab5, N, m, k0, kcount, st, cin, cout, inj);
}

}  // namespace b200
