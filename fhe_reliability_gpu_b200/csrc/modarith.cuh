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
    u32 mprec;    // floor(2^(bits+31)/q): normalised 32-bit reciprocal (big moduli)
    u32 big;      // 1 if q >= 2^36 (32-bit quotient estimate valid for all 64-bit v)
    u32 bits;     // bit length of q
    u32 pshift;   // bits - 33: floor(v/q) ~ umulhi(v >> 32, mprec) >> pshift
    u64 q3;       // 3q, kept in memory so ptxas cannot rematerialise it as an IMAD.WIDE (multiplier pipe)
    u32 delta;    // 2^bits - q when that fits 31 bits (pseudo-Mersenne shape of SEAL's top primes)
    u32 near;     // 1 if the NEAR fast paths are valid: q = 2^bits - delta, delta * 2^(64-bits) < q/2
};

__device__ __forceinline__ u64 pack64(u32 lo, u32 hi) {
    u64 r;
    asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "r"(lo), "r"(hi));
    return r;
}
__device__ __forceinline__ void unpack64(u64 v, u32& lo, u32& hi) {
    asm("mov.b64 {%0,%1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v));
}

// y arbitrary 64-bit, w < q, ws = floor(w 2^64 / q), nq = 2^64 - q  ->  y*w mod q + {0,q}.
// The integer multiplier is the binding pipe of the NTT (profiles/microbench/):
// the quotient is the exact high word (4 IMAD.WIDE), and the remainder
// lo64(y*w) + lo64(k*nq) is ONE multiply-add chain (2 IMAD.WIDE + 4 IMAD) with
// no separate subtraction -- 10 multiplier-pipe instructions, the minimum for an
// exact 64-bit Shoup product on 32-bit multipliers.
__device__ __forceinline__ u64 shoup_mul(u64 y, u64 w, u64 ws, u64 nq) {
    const u64 k = __umul64hi(y, ws);
    u32 yl, yh, wl, wh, kl, kh, nl, nh, rl, rh;
    unpack64(y, yl, yh);
    unpack64(w, wl, wh);
    unpack64(k, kl, kh);
    unpack64(nq, nl, nh);
    u64 r;
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(r) : "r"(yl), "r"(wl));
    asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(r) : "r"(kl), "r"(nl));
    unpack64(r, rl, rh);
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(rh) : "r"(yl), "r"(wh));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(rh) : "r"(yh), "r"(wl));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(rh) : "r"(kl), "r"(nh));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(rh) : "r"(kh), "r"(nl));
    return pack64(rl, rh);
}

// Lazy Shoup product: y arbitrary 64-bit, w < q  ->  y*w mod q + {0,q,2q}  (< 3q).
// The quotient estimate drops only the yl*sl partial product: k = yh*sh + ((yl*sh + yh*sl) >> 32), with the 65-bit cross
// sum kept exactly by a PTX carry chain (mad.lo.cc / madc.hi.cc).  ptxas turns the chain into
//     IMAD.WIDE (yl*sh)  ->  IMAD.HI + carry-out (upper word of yh*sl + that)  ->  SEL (the carry as the pair's high word)
//     ->  IMAD.WIDE (yh*sh + pair)
// i.e. 2 IMAD.WIDE + 1 IMAD.HI + 1 SEL for the quotient and NO separate add: the whole butterfly is 4 IMAD.WIDE + 1 IMAD.HI
// + 4 IMAD + 1 SEL + 4 IADD3 = 14 instructions (the r1 form, which passed zero-extended 32-bit addends to mad.wide, cost
// 5 IMAD.WIDE + 4 IMAD + ~9 IADD3: ptxas splits such addends off into IADD3 pairs).  Multiplier-pipe cycles are about the
// same (~27 against 28); issue slots and ALU work drop by a fifth (profiles/microbench/butterfly_fp64.cu V7).
// Error: k_real - k in [0, 2)  =>  r = y*w - k*q in [0, 2q + q*y/2^64) < 3q for any 64-bit y.
//
// MODE 2 (NEAR: q = 2^bits - delta with delta < 2^31, the shape of every prime CoeffModulus::Create hands out):
// 2^64 - q = (2^32 - 2^sh) 2^32 + delta with sh = bits - 32, so the kl * hi32(-q) term of the remainder is just
// -(kl << sh): one multiplier-pipe IMAD becomes a shift on the ALU pipe.
template <int MODE = 1>
__device__ __forceinline__ u64 shoup_mul_lazy(u64 y, u64 w, u64 ws, u64 nq, u32 nsh = 0) {
    u32 yl, yh, wl, wh, sl, sh, nl, nh, kl, kh, rl, rh;
    unpack64(y, yl, yh);
    unpack64(w, wl, wh);
    unpack64(ws, sl, sh);
    unpack64(nq, nl, nh);
    u64 r;
    asm("{\n\t.reg .u32 t0, t1, t2;\n\t"
        "mul.lo.u32 t0, %2, %5;\n\t"             // lo(yl*sh)
        "mul.hi.u32 t1, %2, %5;\n\t"             // hi(yl*sh)
        "mad.lo.cc.u32 t0, %3, %4, t0;\n\t"      // + lo(yh*sl)
        "madc.hi.cc.u32 t1, %3, %4, t1;\n\t"     // + hi(yh*sl) + carry
        "addc.u32 t2, 0, 0;\n\t"                 // bit 64 of the cross sum
        "mad.lo.cc.u32 %0, %3, %5, t1;\n\t"      // kl = lo(yh*sh) + t1
        "madc.hi.u32 %1, %3, %5, t2;\n\t"        // kh = hi(yh*sh) + t2 + carry
        "}" : "=r"(kl), "=r"(kh) : "r"(yl), "r"(yh), "r"(sl), "r"(sh));
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(r) : "r"(yl), "r"(wl));
    asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(r) : "r"(kl), "r"(nl));
    unpack64(r, rl, rh);
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(rh) : "r"(yl), "r"(wh));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(rh) : "r"(yh), "r"(wl));
    if (MODE == 2) rh -= kl << nsh;
    else asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(rh) : "r"(kl), "r"(nh));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(rh) : "r"(kh), "r"(nl));
    return pack64(rl, rh);
}

// v (any 64-bit) -> congruent value < 2q when BIG (q >= 2^36: 32-bit normalised
// reciprocal, quotient estimate off by at most one), < 3q otherwise.
// NEAR: v = k 2^bits + r  ->  r + k delta  (= v - k q), one IMAD.WIDE; < 2^bits + 2^(64-bits) delta < 2q.
template <int BIG>
__device__ __forceinline__ u64 reduce_lazy(u64 v, const B200Mod& m) {
#ifdef B200_ABL_NOREDUCE     // timing ablation only (wrong results)
    return v;
#endif
    if (BIG == 2) {
        u32 vl, vh;
        unpack64(v, vl, vh);
        const u32 sh = m.bits - 32;
        const u32 k = vh >> sh;
        const u64 low = pack64(vl, vh & ((1u << sh) - 1u));
        u64 r;
        asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(r) : "r"(k), "r"(m.delta), "l"(low));
        return r;
    } else if (BIG) {
        u32 vl, vh, nl, nh, rl, rh;
        unpack64(v, vl, vh);
        unpack64(0 - m.q, nl, nh);
        const u32 k = __umulhi(vh, m.mprec) >> m.pshift;   // floor(v/q) - {0,1}
        u64 r;
        asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(r) : "r"(k), "r"(nl), "l"(v));
        unpack64(r, rl, rh);
        asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(rh) : "r"(k), "r"(nh));
        (void)vl;
        return pack64(rl, rh);
    } else {
        u64 k = __umul64hi(v, m.bmu);
        return v - k * m.q;
    }
}
// bound (in units of q) of reduce_lazy's result
template <int BIG> struct LazyBound { static constexpr int value = BIG ? 2 : 3; };

// v < 3q -> [0,q)
__device__ __forceinline__ u64 canon3(u64 v, u64 q) {
#ifdef B200_ABL_NOCANON
    return v;
#endif
    u64 two_q = q << 1;
    if (v >= two_q) v -= two_q;
    if (v >= q) v -= q;
    return v;
}
// v < 2q -> [0,q)
__device__ __forceinline__ u64 canon2(u64 v, u64 q) {
#ifdef B200_ABL_NOCANON
    return v;
#endif
    return v >= q ? v - q : v;
}

// full a*b mod q, a,b < 2^64 with a*b < q*2^64 is NOT required: generic 128-bit
// Barrett with the two-word ratio (SEAL barrett_reduce_128).  Result in [0,q).
__device__ __forceinline__ u64 barrett128_raw(u64 lo, u64 hi, u64 q, u64 r0, u64 r1) {
    // floor(x * ratio / 2^128), x = hi:lo
    u64 carry = __umul64hi(lo, r0);
    u64 t1lo = lo * r1, t1hi = __umul64hi(lo, r1);
    u64 s = t1lo + carry;
    u64 c1 = t1hi + (s < carry);
    u64 t2lo = hi * r0, t2hi = __umul64hi(hi, r0);
    u64 s2 = s + t2lo;
    u64 c2 = t2hi + (s2 < t2lo);
    u64 quo = hi * r1 + c1 + c2;
    u64 r = lo - quo * q;
    return r >= q ? r - q : r;
}
__device__ __forceinline__ u64 barrett128(u64 lo, u64 hi, const B200Mod& m) {
    return barrett128_raw(lo, hi, m.q, m.ratio0, m.ratio1);
}
__device__ __forceinline__ u64 mulmod_barrett(u64 a, u64 b, const B200Mod& m) {
    return barrett128(a * b, __umul64hi(a, b), m);
}

// (hi:lo) < 2^(2 bits + 4) modulo a pseudo-Mersenne q = 2^bits - delta (50 <= bits <= 60, delta < 2^31: the shape of every prime
// CoeffModulus::Create hands out) by folding 2^bits = delta three times -- 5 IMAD.WIDE against the ~17 of barrett128:
//   V1 = (V  >> bits) delta + (V  mod 2^bits) < 2^(bits + 35) + 2^bits
//   V2 = (V1 >> bits) delta + (V1 mod 2^bits) < 2^66 + 2^bits
//   V3 = (V2 >> bits) delta + (V2 mod 2^bits) < 2^(97 - bits) + 2^bits < 2 q      -> one conditional subtraction.  Result in [0,q).
// One long dependent chain: pays in throughput-bound streaming kernels, not in latency-bound ones (profiles/r2r_tc5_nearfold.txt).
__device__ __forceinline__ u64 fold128_near(u64 lo, u64 hi, u64 q, u32 delta, u32 bits) {
    const u64 mask = ((u64)1 << bits) - 1;
    const u64 h1 = (lo >> bits) | (hi << (64 - bits));
    u64 p0, p1;
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(p0) : "r"((u32)h1), "r"(delta));
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(p1) : "r"((u32)(h1 >> 32)), "r"(delta));
    u64 v1lo = p0 + (p1 << 32);
    u64 v1hi = (p1 >> 32) + (v1lo < p0);
    const u64 l1 = lo & mask;
    v1lo += l1;
    v1hi += (v1lo < l1);
    const u64 h2 = (v1lo >> bits) | (v1hi << (64 - bits));
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(p0) : "r"((u32)h2), "r"(delta));
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(p1) : "r"((u32)(h2 >> 32)), "r"(delta));
    u64 v2lo = p0 + (p1 << 32);
    u64 v2hi = (p1 >> 32) + (v2lo < p0);
    const u64 l2 = v1lo & mask;
    v2lo += l2;
    v2hi += (v2lo < l2);
    const u32 h3 = (u32)((v2lo >> bits) | (v2hi << (64 - bits)));
    u64 v3;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(v3) : "r"(h3), "r"(delta), "l"(v2lo & mask));
    return v3 >= q ? v3 - q : v3;
}
// (hi:lo) mod q: the fold when the modulus has the shape and the value is in its range, the general Barrett otherwise
__device__ __forceinline__ u64 reduce128(u64 lo, u64 hi, const B200Mod& m) {
    if (m.near && m.bits >= 50 && (hi >> (2 * m.bits - 60)) == 0) return fold128_near(lo, hi, m.q, m.delta, m.bits);
    return barrett128(lo, hi, m);
}

__device__ __forceinline__ ulonglong2 ldg_tw(const ulonglong2* p) {
    return __ldg(p);
}
