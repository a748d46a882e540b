#include "numth.hpp"
#include <map>

namespace b200 {
namespace nt {

bool is_prime(uint64_t n) {
    // deterministic Miller-Rabin, valid for all 64-bit n
    static const uint64_t bases[12] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    if (n < 2) return false;
    for (uint64_t b : bases) {
        if (n == b) return true;
        if (n % b == 0) return false;
    }
    uint64_t d = n - 1;
    int r = 0;
    while (!(d & 1)) { d >>= 1; r++; }
    for (uint64_t b : bases) {
        uint64_t x = powmod(b, d, n);
        if (x == 1 || x == n - 1) continue;
        bool comp = true;
        for (int j = 1; j < r && comp; j++) {
            x = mulmod(x, x, n);
            if (x == n - 1) comp = false;
        }
        if (comp) return false;
    }
    return true;
}

bool primes_create(uint64_t N, const int* bits, int n, uint64_t* out) {
    // SEAL: per bit size, walk down from the largest value < 2^bits that is
    // 1 mod 2N; Create() then pops from the back (smallest of the top-k first).
    std::map<int, int> need;
    for (int i = 0; i < n; i++) {
        if (bits[i] < 2 || bits[i] > 60) return false;
        need[bits[i]]++;
    }
    std::map<int, std::vector<uint64_t>> tab;
    const uint64_t factor = 2 * N;
    for (auto& kv : need) {
        const int b = kv.first;
        uint64_t value = (((uint64_t)1 << b) - 1) / factor * factor + 1;
        const uint64_t lower = (uint64_t)1 << (b - 1);
        auto& v = tab[b];
        while ((int)v.size() < kv.second && value > lower) {
            if (is_prime(value)) v.push_back(value);
            if (value < factor) break;
            value -= factor;
        }
        if ((int)v.size() < kv.second) return false;
    }
    for (int i = 0; i < n; i++) {
        auto& v = tab[bits[i]];
        out[i] = v.back();
        v.pop_back();
    }
    return true;
}

uint64_t min_primitive_root(uint64_t two_n, uint64_t q) {
    if (two_n < 2 || (q - 1) % two_n) return 0;
    const uint64_t e = (q - 1) / two_n;
    uint64_t root = 0;
    for (uint64_t g = 2; g < q && g < 100000; g++) {
        uint64_t r = powmod(g, e, q);
        if (powmod(r, two_n >> 1, q) == q - 1) { root = r; break; }   // order exactly 2N
    }
    if (!root) return 0;
    // all primitive 2N-th roots are the odd powers of `root`
    const uint64_t sq = mulmod(root, root, q);
    uint64_t cur = root, best = root;
    for (uint64_t i = 0; i < two_n / 2; i++) {
        if (cur < best) best = cur;
        cur = mulmod(cur, sq, q);
    }
    return best;
}

void const_ratio(uint64_t q, uint64_t ratio[2]) {
    u128 all = ~(u128)0;
    u128 quo = all / q;
    if (all % q + 1 == q) quo += 1;
    ratio[0] = (uint64_t)quo;
    ratio[1] = (uint64_t)(quo >> 64);
}

void tables(int log_n, uint64_t q, uint64_t psi, uint64_t* rp, uint64_t* rps, uint64_t* irp, uint64_t* irps) {
    const uint64_t N = (uint64_t)1 << log_n, ipsi = invmod(psi, q);
    uint64_t p = 1, ip = 1;
    for (uint64_t i = 0; i < N; i++) {
        const uint64_t br = bitrev(i, log_n);
        if (rp) rp[br] = p;
        if (rps) rps[br] = shoup(p, q);
        if (irp) irp[br] = ip;
        if (irps) irps[br] = shoup(ip, q);
        p = mulmod(p, psi, q);
        ip = mulmod(ip, ipsi, q);
    }
}

void forward_ntt_host(uint64_t* a, int log_n, uint64_t q, const uint64_t* rp) {
    const uint64_t N = (uint64_t)1 << log_n;
    uint64_t t = N >> 1;
    for (uint64_t m = 1; m < N; m <<= 1, t >>= 1)
        for (uint64_t i = 0; i < m; i++) {
            const uint64_t w = rp[m + i];
            uint64_t* x = a + 2 * i * t;
            for (uint64_t j = 0; j < t; j++) {
                const uint64_t u = x[j], v = mulmod(x[j + t], w, q);
                x[j] = u + v >= q ? u + v - q : u + v;
                x[j + t] = u >= v ? u - v : u + q - v;
            }
        }
}

void abft_output_weights(int log_n, uint64_t q, uint64_t psi, uint64_t* omega) {
    const uint64_t N = (uint64_t)1 << log_n;
    std::vector<uint64_t> irp(N);
    tables(log_n, q, psi, nullptr, nullptr, irp.data(), nullptr);
    for (uint64_t i = 0; i < N; i++) omega[i] = abft_input_weight(i, log_n) % q;
    forward_ntt_host(omega, log_n, q, irp.data());
    const uint64_t ninv = invmod(N % q, q);
    for (uint64_t i = 0; i < N; i++) omega[i] = mulmod(omega[i], ninv, q);
}

}  // namespace nt
}  // namespace b200
