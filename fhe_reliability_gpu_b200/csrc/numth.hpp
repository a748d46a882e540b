// Host number theory for context construction (product code; independent of
// oracle/).  Restates the rules the reference harness relies on from
// SEAL/Phantom: CoeffModulus::Create prime selection (pinned by
// reliability_test/data/bits1-16_num1.txt:10), minimal primitive 2N-th root,
// bit-reversed root-power tables with Shoup quotients (ntt_test.cu:60-68) and
// the two-word Barrett ratio (ntt_test.cu:49-53).
#pragma once
#include <cstdint>
#include <vector>

namespace b200 {
namespace nt {

typedef unsigned __int128 u128;

inline uint64_t mulmod(uint64_t a, uint64_t b, uint64_t q) { return (uint64_t)((u128)a * b % q); }
inline uint64_t powmod(uint64_t b, uint64_t e, uint64_t q) {
    uint64_t r = 1 % q;
    b %= q;
    for (; e; e >>= 1) {
        if (e & 1) r = mulmod(r, b, q);
        b = mulmod(b, b, q);
    }
    return r;
}
inline uint64_t invmod(uint64_t a, uint64_t q) { return powmod(a, q - 2, q); }
inline uint64_t shoup(uint64_t w, uint64_t q) { return (uint64_t)(((u128)w << 64) / q); }
inline uint64_t bitrev(uint64_t x, int bits) {
    uint64_t r = 0;
    for (int i = 0; i < bits; i++, x >>= 1) r = (r << 1) | (x & 1);
    return r;
}
inline int bitlen(uint64_t x) {
    int b = 0;
    for (; x; x >>= 1) b++;
    return b;
}

bool is_prime(uint64_t n);
// returns false if not enough primes of some size exist
bool primes_create(uint64_t N, const int* bits, int n, uint64_t* out);
// 0 if q != 1 mod 2N or no root found
uint64_t min_primitive_root(uint64_t two_n, uint64_t q);
void const_ratio(uint64_t q, uint64_t ratio[2]);
// rp/rps/irp/irps may be null
void tables(int log_n, uint64_t q, uint64_t psi, uint64_t* rp, uint64_t* rps, uint64_t* irp, uint64_t* irps);

// SEAL-order forward negacyclic NTT on the host (table construction only, O(N log N) 128-bit mulmods)
void forward_ntt_host(uint64_t* a, int log_n, uint64_t q, const uint64_t* rp);

// ABFT weighted checksum (rfhe_framewk/src/negaclic_ntt.py:130-149 ecc_check): input weights are
// generate_weights(N) (:8-13): w_i = (i mod p) + 1 + (i div p) + 1 with period p = sqrt(N)
// (p = 2^ceil(log_n/2) when N is not a perfect square).  Small (< 2^10), so the device
// accumulates sum w_i x_i with two carry-free IMAD.WIDE per coefficient.
inline int abft_period_log(int log_n) { return (log_n + 1) / 2; }
inline uint64_t abft_input_weight(uint64_t i, int log_n) {
    const int pl = abft_period_log(log_n);
    return (i & (((uint64_t)1 << pl) - 1)) + (i >> pl) + 2;
}
// omega with  sum_k omega_k NTT(x)_k == sum_i w_i x_i (mod q) for every x, NTT in SEAL (bit-reversed)
// output order:  omega = N^-1 * NTT_{psi^-1}(w)  (the prototype's w_hat = INTT(D^-1 w), :137-140)
void abft_output_weights(int log_n, uint64_t q, uint64_t psi, uint64_t* omega);

}  // namespace nt
}  // namespace b200
