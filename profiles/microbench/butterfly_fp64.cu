// Register-only radix-16 Cooley-Tukey phase throughput on sm_100a (VERDICT r1 "next" item 4): can the FP64 pipe,
// idle in every NTT kernel, take the quotient estimate of the lazy Shoup product off the integer multiplier?
//
//   V2  the r1 product: k = yh sh + hi32(yl sh) + hi32(yh sl + lo32(yl sh))      5 IMAD.WIDE + 4 IMAD, r < 3q
//   V3  integer, cheaper estimate: k = yh sh + hi32(yl sh) + hi32(yh sl)         3 IMAD.WIDE + 2 IMAD.HI + 4 IMAD, r < 4q
//   V7  integer, the cross sum kept exactly by a PTX carry chain (mad.lo.cc / madc.hi.cc): ptxas emits IMAD.WIDE, IMAD.HI with
//       carry-out, SEL, IMAD.WIDE and no separate adds                             4 IMAD.WIDE + 1 IMAD.HI + 4 IMAD + 1 SEL, r < 3q
//   V6  FP64 estimate: k = yh sh + int(fma(Yl, a, fma(Yh, b, C)))                3 IMAD.WIDE + 4 IMAD + 2 DFMA,    r < 4q
//
// V6: k = floor(y ws / 2^64) = yh sh + (yl sh + yh sl) / 2^32 + (< 1).  The middle term S is < 2^33 and only its integer part
// matters, so it is formed by two DFMAs on exactly representable operands, with NO conversion instruction:
//     a = sh 2^-32, b = sl 2^-32, C = M - 1 - (sh + sl) 2^20, M = 2^52 + 2^34              (per twiddle, prepared)
//     Yl = 2^52 + yl, Yh = 2^52 + yh    = the register pairs {yl, 0x43300000}, {yh, 0x43300000}
//     t  = fma(Yl, a, fma(Yh, b, C)) = M - 1 + S + e,  |e| <= 1 (two roundings at magnitudes < 2^53, ulp <= 1)
// and the integer sits in t's low mantissa bits: k' = yh sh + bits(t) = k_est + (0x43300004 << 32), the constant is taken out
// of the product's high word by the butterfly's own 3-input adds.  k_est - k_real in (-3, 0]  =>  r = y w - k_est q in [0, 4q)
// for ANY 64-bit y (CPU emulation, 1.2e8 cases: max 3.67 q).  To avoid building {y?, 0x43300000} pairs with MOVs in every
// butterfly (ptxas emits 4 per butterfly, which eats the gain), a coefficient is HELD as those two pairs: the 32-bit adds of the
// butterfly write the low registers, the high registers keep 0x43300000 (4 registers per coefficient instead of 2).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o butterfly_fp64 butterfly_fp64.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
typedef unsigned long long u64;
typedef unsigned int u32;

__device__ __forceinline__ u64 pack64(u32 lo, u32 hi) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "r"(lo), "r"(hi)); return r; }
__device__ __forceinline__ void unpack64(u64 v, u32& lo, u32& hi) { asm("mov.b64 {%0,%1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v)); }

struct ModC {
    u64 q, nq, q3, q4, q4p;   // q4p = 4q + (Kc << 32)
    u32 negKc;                // -(0x43300004 * lo32(nq))
    u32 sh, delta;            // NEAR reduce: q = 2^(32+sh) - delta
};
struct TwRaw { u64 w, ws; };                        // 16 bytes: what the tables hold today
struct TwF { u64 w; u32 sh, pad; double a, b, C; }; // 40 bytes: prepared form of V6

// ---- V2 / V3: plain u64 coefficients -------------------------------------------------------------------------
template <int V>
__device__ __forceinline__ void bf_int(u64& X, u64& Y, const TwRaw& t, const ModC& m) {
    u32 yl, yh, wl, wh, sl, sh, nl, nh, kl, kh, rl, rh;
    unpack64(Y, yl, yh); unpack64(t.w, wl, wh); unpack64(t.ws, sl, sh); unpack64(m.nq, nl, nh);
    u64 k, r;
    if (V == 2) {
        u32 c1l, c1h;
        u64 c1, s2;
        asm("mul.wide.u32 %0, %1, %2;" : "=l"(c1) : "r"(yl), "r"(sh));
        unpack64(c1, c1l, c1h);
        asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(s2) : "r"(yh), "r"(sl), "l"((u64)c1l));
        asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(k) : "r"(yh), "r"(sh), "l"((u64)c1h));
        k += (s2 >> 32);
    } else if (V == 7) {   // carry chain: k = yh sh + ((yl sh + yh sl) >> 32) exactly; ptxas: WIDE, IMAD.HI + carry-out, SEL, WIDE
        asm("{\n\t.reg .u32 t0, t1, t2;\n\t"
            "mul.lo.u32 t0, %2, %5;\n\t"
            "mul.hi.u32 t1, %2, %5;\n\t"
            "mad.lo.cc.u32 t0, %3, %4, t0;\n\t"
            "madc.hi.cc.u32 t1, %3, %4, t1;\n\t"
            "addc.u32 t2, 0, 0;\n\t"
            "mad.lo.cc.u32 %0, %3, %5, t1;\n\t"
            "madc.hi.u32 %1, %3, %5, t2;\n\t"
            "}" : "=r"(kl), "=r"(kh) : "r"(yl), "r"(yh), "r"(sl), "r"(sh));
        k = pack64(kl, kh);
    } else {
        u32 h1, h2;
        asm("mul.hi.u32 %0, %1, %2;" : "=r"(h1) : "r"(yl), "r"(sh));
        asm("mul.hi.u32 %0, %1, %2;" : "=r"(h2) : "r"(yh), "r"(sl));
        asm("mul.wide.u32 %0, %1, %2;" : "=l"(k) : "r"(yh), "r"(sh));
        k = k + h1 + h2;
    }
    unpack64(k, kl, kh);
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(r) : "r"(yl), "r"(wl));
    asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(r) : "r"(kl), "r"(nl));
    unpack64(r, rl, rh);
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(rh) : "r"(yl), "r"(wh));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(rh) : "r"(yh), "r"(wl));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(rh) : "r"(kl), "r"(nh));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(rh) : "r"(kh), "r"(nl));
    const u64 T = pack64(rl, rh), x = X;
    X = x + T;
    Y = x - T + (V == 2 || V == 7 ? m.q3 : m.q4);            // V3: k_est in (k_real - 3, k_real] -> r < 3q + q y / 2^64 < 4q
}
__device__ __forceinline__ u64 reduce_near(u64 v, const ModC& m) {   // any v -> < 2q: one IMAD.WIDE
    u32 vl, vh;
    unpack64(v, vl, vh);
    const u32 k = vh >> m.sh;
    u64 r;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(r) : "r"(k), "r"(m.delta), "l"(pack64(vl, vh & ((1u << m.sh) - 1u))));
    return r;
}

// ---- V6: coefficients held as two magic doubles ---------------------------------------------------------------
struct Co { double l, h; };    // {lo, 0x43300000}, {hi, 0x43300000}
__device__ __forceinline__ u32 lo32(double d) { return (u32)__double2loint(d); }
__device__ __forceinline__ double setlo(double d, u32 lo) { return __hiloint2double(__double2hiint(d), (int)lo); }
__device__ __forceinline__ TwF tw_prep(u64 w, u64 ws) {
    u32 sl, sh;
    unpack64(ws, sl, sh);
    TwF t;
    t.w = w; t.sh = sh; t.pad = 0;
    t.a = fma(__hiloint2double(0x43300000, sh), 0x1p-32, -0x1p20);
    t.b = fma(__hiloint2double(0x43300000, sl), 0x1p-32, -0x1p20);
    t.C = fma(-0x1p52, t.b, fma(-0x1p52, t.a, 0x1p52 + 0x1p34 - 1.0));
    return t;
}
__device__ __forceinline__ void bf_f64(Co& A, Co& B, const TwF& t, const ModC& m) {
    const u32 yl = lo32(B.l), yh = lo32(B.h);
    u32 wl, wh, nl, nh, kl, kh, rl, rh;
    unpack64(t.w, wl, wh); unpack64(m.nq, nl, nh);
    double d = fma(B.h, t.b, t.C);
    d = fma(B.l, t.a, d);
    u64 k, r;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(k) : "r"(yh), "r"(t.sh), "l"((u64)__double_as_longlong(d)));
    unpack64(k, kl, kh);
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(r) : "r"(yl), "r"(wl));
    asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(r) : "r"(kl), "r"(nl));
    unpack64(r, rl, rh);
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(rh) : "r"(yl), "r"(wh));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(rh) : "r"(yh), "r"(wl));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(rh) : "r"(kl), "r"(nh));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(rh) : "r"(kh), "r"(nl));
    const u32 xl = lo32(A.l), xh = lo32(A.h);
    u32 al, ah, bl, bh;
    asm("add.cc.u32 %0, %2, %4;\n\taddc.u32 %1, %3, %5;" : "=r"(al), "=r"(ah) : "r"(xl), "r"(xh), "r"(rl), "r"(rh));
    ah += m.negKc;                                   // folds into the IADD3.X (3 inputs)
    const u64 bn = pack64(xl, xh) - pack64(rl, rh) + m.q4p;
    unpack64(bn, bl, bh);
    A.l = setlo(A.l, al); A.h = setlo(A.h, ah);
    B.l = setlo(B.l, bl); B.h = setlo(B.h, bh);
}
__device__ __forceinline__ void reduce_near(Co& c, const ModC& m) {
    u32 rl, rh;
    unpack64(reduce_near(pack64(lo32(c.l), lo32(c.h)), m), rl, rh);
    c.l = setlo(c.l, rl); c.h = setlo(c.h, rh);
}

#define ITERS 32
// One radix-16 phase per iteration: 32 butterflies, 15 twiddles from shared memory.
// TWM 0: one list shared by the CTA (column passes; V6 reads the prepared 40-byte form)
// TWM 1: one list per thread, raw (w, ws) entries, stride = blockDim (row passes; V6 prepares each twiddle on the fly)
template <int V, int TWM>
__global__ void __launch_bounds__(256, V == 6 ? 1 : 2) k(u64* out, const ModC* mp, const TwRaw* twg, long long* cyc, u32 K, int check) {
    extern __shared__ __align__(16) unsigned char smraw[];
    TwRaw* sraw = reinterpret_cast<TwRaw*>(smraw);
    TwF* sprep = reinterpret_cast<TwF*>(smraw);
    if (TWM == 0) {
        if (threadIdx.x < 15) {
            const TwRaw t = twg[(threadIdx.x * 977 + blockIdx.x) & 4095];
            if (V == 6) sprep[threadIdx.x] = tw_prep(t.w, t.ws);
            else sraw[threadIdx.x] = t;
        }
    } else {
        for (int j = 0; j < 15; j++) sraw[j * blockDim.x + threadIdx.x] = twg[(j * 977 + threadIdx.x * 31 + blockIdx.x) & 4095];
    }
    __syncthreads();
    const ModC m = *mp;
    u64 v[16];
    Co c[16];
#pragma unroll
    for (int e = 0; e < 16; e++) {
        const u64 x = (0x9E3779B97F4A7C15ull * (e + 1) + threadIdx.x * 0x100000001B3ull + blockIdx.x) >> 5;
        v[e] = x;
        c[e].l = __hiloint2double(K, (u32)x);
        c[e].h = __hiloint2double(K, (u32)(x >> 32));
    }
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int e = 0; e < 8; e++) {        // phase start: X half below 2q (what reduce_lazy does in the kernels)
            if (V == 6) reduce_near(c[e], m); else v[e] = reduce_near(v[e], m);
        }
#pragma unroll
        for (int t = 0; t < 4; t++) {
            const int half = 8 >> t;
            if (V != 2 && V != 7 && t == 3) {          // products < 4q: X path 2 + 3*4 = 14q, one more sub-stage would pass 16q
#pragma unroll
                for (int e = 0; e < 16; e += 2) { if (V == 6) reduce_near(c[e], m); else v[e] = reduce_near(v[e], m); }
            }
#pragma unroll
            for (int g = 0; g < (1 << t); g++) {
                const int j = (1 << t) - 1 + g;
                if (V == 6) {
                    TwF tw;
                    if (TWM == 0) tw = sprep[j];
                    else { const TwRaw r = sraw[j * blockDim.x + threadIdx.x]; tw = tw_prep(r.w, r.ws); }
#pragma unroll
                    for (int i = 0; i < half; i++) bf_f64(c[g * 2 * half + i], c[g * 2 * half + i + half], tw, m);
                } else {
                    const TwRaw tw = TWM == 0 ? sraw[j] : sraw[j * blockDim.x + threadIdx.x];
#pragma unroll
                    for (int i = 0; i < half; i++) bf_int<V>(v[g * 2 * half + i], v[g * 2 * half + i + half], tw, m);
                }
            }
        }
    }
    long long t1 = clock64();
    const size_t o = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 16;
    if (check) {
#pragma unroll
        for (int e = 0; e < 16; e++) out[o + e] = V == 6 ? pack64(lo32(c[e].l), lo32(c[e].h)) : v[e];
    } else {
        u64 acc = 0;
#pragma unroll
        for (int e = 0; e < 16; e++) acc ^= V == 6 ? pack64(lo32(c[e].l), lo32(c[e].h)) : v[e];
        out[o] = acc;
    }
    if (threadIdx.x == 0) atomicMax((unsigned long long*)cyc, (unsigned long long)(t1 - t0));
}

static u64 mulmod(u64 a, u64 b, u64 q) { return (u64)((unsigned __int128)a * b % q); }

template <int V, int TWM>
void run(const char* name, const TwRaw* d_tw, const TwRaw* h_tw, const ModC* d_m, u64 q) {
    u64* out;
    long long* cyc;
    cudaMalloc(&out, (size_t)148 * 8 * 256 * 16 * 8);
    cudaMallocManaged(&cyc, 8);
    const size_t smem_per_thread = TWM == 0 ? 0 : 15 * 16, smem_fixed = TWM == 0 ? 15 * sizeof(TwF) : 0;
    cudaFuncSetAttribute(k<V, TWM>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    {   // correctness: residues after ITERS phases against exact CPU arithmetic
        const int threads = 64;
        k<V, TWM><<<2, threads, smem_fixed + smem_per_thread * threads>>>(out, d_m, d_tw, cyc, 0x43300000u, 1);
        cudaDeviceSynchronize();
        u64* h = (u64*)malloc((size_t)2 * threads * 16 * 8);
        cudaMemcpy(h, out, (size_t)2 * threads * 16 * 8, cudaMemcpyDeviceToHost);
        long bad = 0;
        for (int b = 0; b < 2; b++)
            for (int th = 0; th < threads; th++) {
                u64 v[16];
                for (int e = 0; e < 16; e++) v[e] = ((0x9E3779B97F4A7C15ull * (e + 1) + th * 0x100000001B3ull + b) >> 5) % q;
                for (int it = 0; it < ITERS; it++)
                    for (int t = 0; t < 4; t++) {
                        const int half = 8 >> t;
                        for (int g = 0; g < (1 << t); g++) {
                            const int j = (1 << t) - 1 + g;
                            const u64 w = TWM == 0 ? h_tw[(j * 977 + b) & 4095].w : h_tw[(j * 977 + th * 31 + b) & 4095].w;
                            for (int i = 0; i < half; i++) {
                                const int a = g * 2 * half + i, c = a + half;
                                const u64 T = mulmod(v[c], w, q), x = v[a];
                                v[a] = (x + T) % q;
                                v[c] = (x + q - T) % q;
                            }
                        }
                    }
                for (int e = 0; e < 16; e++) bad += h[((size_t)b * threads + th) * 16 + e] % q != v[e];
            }
        printf("%-46s residues vs exact: %s (%ld mismatches)\n", name, bad ? "FAIL" : "ok", bad);
        free(h);
    }
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, k<V, TWM>);
    for (int threads = 128; threads <= 256; threads += 128) {
        int occ = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k<V, TWM>, threads, smem_fixed + smem_per_thread * threads);
        for (int ctas = 1; ctas <= occ && ctas <= 8; ctas++) {      // ctas CTAs resident per SM, one wave
            if (threads * ctas < 256 || (ctas > 2 && (ctas & 1))) continue;
            for (int rep = 0; rep < 2; rep++) {
                *cyc = 0;
                k<V, TWM><<<148 * ctas, threads, smem_fixed + smem_per_thread * threads>>>(out, d_m, d_tw, cyc, 0x43300000u, 0);
                cudaDeviceSynchronize();
            }
            const double warp_bf_per_smsp = (double)ITERS * 32 * (threads / 32) * ctas / 4.0;
            printf("%-46s regs=%3d warps/SM=%2d  cycles=%8lld  cyc/warp-butterfly/SMSP=%.2f\n", name, fa.numRegs, threads / 32 * ctas, *cyc,
                   *cyc / warp_bf_per_smsp);
        }
    }
    cudaFree(out);
    cudaFree(cyc);
}

int main() {
    const u64 q = 1152921504606584833ull;     // a NEAR modulus of CoeffModulus::Create(2^16, {60})
    TwRaw* h_tw = (TwRaw*)malloc(4096 * sizeof(TwRaw));
    u64 s = 88172645463325252ull;
    for (int i = 0; i < 4096; i++) {
        s ^= s << 13; s ^= s >> 7; s ^= s << 17;
        const u64 w = s % q;
        h_tw[i].w = w;
        h_tw[i].ws = (u64)((((unsigned __int128)w) << 64) / q);
    }
    TwRaw* d_tw;
    cudaMalloc(&d_tw, 4096 * sizeof(TwRaw));
    cudaMemcpy(d_tw, h_tw, 4096 * sizeof(TwRaw), cudaMemcpyHostToDevice);
    ModC m;
    m.q = q; m.nq = 0 - q; m.q3 = 3 * q; m.q4 = 4 * q;
    const u32 Kc = 0x43300004u * (u32)m.nq;
    m.negKc = 0u - Kc;
    m.q4p = m.q4 + ((u64)Kc << 32);
    m.sh = 60 - 32; m.delta = (u32)((1ull << 60) - q);
    ModC* d_m;
    cudaMalloc(&d_m, sizeof(ModC));
    cudaMemcpy(d_m, &m, sizeof(ModC), cudaMemcpyHostToDevice);
    run<2, 0>("V2 int 5W+4I, CTA-shared twiddles", d_tw, h_tw, d_m, q);
    run<7, 0>("V7 int carry chain 4W+1HI+4I, CTA-shared", d_tw, h_tw, d_m, q);
    run<7, 1>("V7 int carry chain 4W+1HI+4I, per-thread", d_tw, h_tw, d_m, q);
    run<3, 0>("V3 int 3W+2HI+4I, CTA-shared twiddles", d_tw, h_tw, d_m, q);
    run<6, 0>("V6 fp64 3W+4I+2DFMA, CTA-shared prepared", d_tw, h_tw, d_m, q);
    run<2, 1>("V2 int 5W+4I, per-thread twiddles", d_tw, h_tw, d_m, q);
    run<3, 1>("V3 int 3W+2HI+4I, per-thread twiddles", d_tw, h_tw, d_m, q);
    run<6, 1>("V6 fp64, per-thread raw twiddles + prep", d_tw, h_tw, d_m, q);
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
