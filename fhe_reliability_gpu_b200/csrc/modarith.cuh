// Device-side modular arithmetic for 64-bit RNS limbs (q < 2^60, SEAL's user
// modulus cap).  Replaces the arithmetic that lives in phantom-fhe's
// uintmodmath.cuh / SEAL's uintarithsmallmod.h (both absent from the reference
// tree); formulas: Shoup multiply with precomputed quotient (the tables that
// reliability_test/ntt_test.cu:63-64 uploads) and Barrett reduction
// (rfhe_framewk/src/barrett_final.cpp:68-79).
//
// Lazy-range discipline (DESIGN.md "bounds"): because q < 2^60, any value
// below 16q fits a word.  A Shoup product of ANY 64-bit y lands in [0,2q), so
// Cooley-Tukey butterflies only grow the X path by 2q per stage and a radix-16
// or radix-32 register phase needs a single cheap quotient-estimate reduction
// (to < 3q) on the X half of its first stage.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

typedef unsigned long long u64;
typedef unsigned int u32;

// Per-modulus constants, one per RNS limb, resident in device memory.
struct __align__(16) B200Mod {
    u64 q;        // modulus
    u64 bmu;      // floor((2^64-1)/q): one-word Barrett ratio (exact-ish reduce)
    u64 ninv;     // N^-1 mod q
    u64 ninv_s;   // Shoup quotient of ninv
    u64 iw1n;     // itw[1] * N^-1 mod q (last inverse stage, SEAL folds the scale)
    u64 iw1n_s;   // Shoup quotient of iw1n
    u64 ratio0;   // floor(2^128/q) low word  (DModulus::set arg 2, ntt_test.cu:49-53)
    u64 ratio1;   // floor(2^128/q) high word (DModulus::set arg 3)
    u32 mu32;     // floor(2^64/q) (fits 32 bits when q >= 2^36): quotient estimate
    u32 big;      // 1 if q >= 2^36 (fast quotient estimate valid for all 64-bit v)
    u32 bits;     // bit length of q
    u32 pad;
};

__device__ __forceinline__ u64 shoup_mul(u64 y, u64 w, u64 ws, u64 q) {
    // y arbitrary 64-bit, w < q, ws = floor(w 2^64 / q)  ->  y*w mod q + {0,q}
    u64 hi = __umul64hi(y, ws);
    return y * w - hi * q;
}

// v (any 64-bit when BIG, else any 64-bit too but slower) -> congruent value < 3q
template <bool BIG>
__device__ __forceinline__ u64 reduce_lazy(u64 v, const B200Mod& m) {
    if (BIG) {
        u32 k = __umulhi((u32)(v >> 32), m.mu32);   // floor(v/q) - {0,1,2}
        return v - (u64)k * m.q;
    } else {
        u64 k = __umul64hi(v, m.bmu);
        return v - k * m.q;
    }
}

// v < 3q -> [0,q)
__device__ __forceinline__ u64 canon3(u64 v, u64 q) {
    u64 two_q = q << 1;
    if (v >= two_q) v -= two_q;
    if (v >= q) v -= q;
    return v;
}
// v < 2q -> [0,q)
__device__ __forceinline__ u64 canon2(u64 v, u64 q) { return v >= q ? v - q : v; }

// full a*b mod q, a,b < 2^64 with a*b < q*2^64 is NOT required: generic 128-bit
// Barrett with the two-word ratio (SEAL barrett_reduce_128).  Result in [0,q).
__device__ __forceinline__ u64 barrett128(u64 lo, u64 hi, const B200Mod& m) {
    // floor(x * ratio / 2^128), x = hi:lo
    u64 r0 = m.ratio0, r1 = m.ratio1;
    u64 carry = __umul64hi(lo, r0);
    u64 t1lo = lo * r1, t1hi = __umul64hi(lo, r1);
    u64 s = t1lo + carry;
    u64 c1 = t1hi + (s < carry);
    u64 t2lo = hi * r0, t2hi = __umul64hi(hi, r0);
    u64 s2 = s + t2lo;
    u64 c2 = t2hi + (s2 < t2lo);
    u64 quo = hi * r1 + c1 + c2;
    u64 r = lo - quo * m.q;
    return r >= m.q ? r - m.q : r;
}
__device__ __forceinline__ u64 mulmod_barrett(u64 a, u64 b, const B200Mod& m) {
    return barrett128(a * b, __umul64hi(a, b), m);
}

__device__ __forceinline__ ulonglong2 ldg_tw(const ulonglong2* p) {
    return __ldg(p);
}
