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

}  // namespace nt
}  // namespace b200
