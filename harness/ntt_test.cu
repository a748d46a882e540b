// ntt_test -- fault-propagation harness of the reliability study, on libb200fhe.so.
//
// Same CLI, stdout and stderr contract as reliability_test/ntt_test.cu:201-220 (CLI),
// :125-127 (flip announcement), :171-198 (report), which the sweep scripts parse
// (test_scripts/gen_errorimpact.py:29-30, run_bench_test.sh:10, run_real_test.sh:25):
//   ntt_test <log_dim> <batch_size> <num_flips> <num_target_symbols> [seed]
// A batch of `batch_size` limbs (one distinct 50-bit prime each, ntt_test.cu:44) is transformed
// clean, then again after XOR-ing `num_flips` distinct bit positions into each of
// `num_target_symbols` distinct words of the input; outputs are compared bit by bit.
// Differences from the reference program, all opt-in: an optional 5th argument (or
// B200FHE_SEED) makes the run reproducible (the reference seeds from std::random_device,
// ntt_test.cu:73-74), and B200FHE_PRIME_BITS selects another prime size (default 50).
// With num_flips == 0 this is ntt_real_test.cu (injection disabled, :108-117).
#include <cuda_runtime.h>
#include <cinttypes>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <random>
#include <set>
#include <string>
#include <vector>
#include "b200fhe.h"

static void die(const char* what) {
    std::fprintf(stderr, "ERROR: %s: %s\n", what, b200fhe_last_error());
    std::exit(2);
}
#define OK(call) do { if ((call) != 0) die(#call); } while (0)

int main(int argc, char** argv) {
    // Two CLIs, both kept: the current one, `<log_dim> <batch_size> <num_flips> <num_target_symbols>`
    // (ntt_test.cu:201-216), and the earlier 3-argument one that reliability_test/run_bench_test.sh:10 still calls and
    // exp_log.txt:3-8 records: `<log_dim> <batch_size> <n>` = flip n unique random bits anywhere in the data buffer.
    if (argc < 4) {
        std::cerr << "Usage: " << argv[0] << " <log_dim> <batch_size> <num_flips> <num_target_symbols>\n"
                  << "   or: " << argv[0] << " <log_dim> <batch_size> <num_unique_bit_flips>\n";
        return 1;
    }
    const bool legacy = argc == 4;
    const size_t log_dim = std::stoul(argv[1]), batch = std::stoul(argv[2]);
    const int num_flips = legacy ? 1 : std::stoi(argv[3]), num_syms = legacy ? 1 : std::stoi(argv[4]);
    const long long legacy_bits = legacy ? std::stoll(argv[3]) : 0;
    const bool real_mode = std::getenv("B200FHE_REAL_TEST") != nullptr;   // ntt_real_test: no injection
    if (!real_mode && (num_flips < 1 || num_syms < 1)) {
        std::cerr << "Error: <num_flips> and <num_target_symbols> must be >= 1\n";
        return 1;
    }
    const size_t dim = (size_t)1 << log_dim;
    const uint64_t total = (uint64_t)batch * dim, total_bits = total * 64;
    if ((uint64_t)num_flips > 64) {
        std::cerr << "ERROR: num_flips (" << num_flips << ") exceeds bits per symbol (64).\n";
        return 1;
    }
    if (legacy && (legacy_bits < 1 || (uint64_t)legacy_bits > total_bits)) {
        std::cerr << "ERROR: number of bit flips must be in [1, " << total_bits << "]\n";
        return 1;
    }
    if ((uint64_t)num_syms > total) {
        std::cerr << "ERROR: num_target_symbols (" << num_syms << ") exceeds total symbols (" << total << ").\n";
        return 1;
    }
    uint64_t seed;
    if (!legacy && argc > 5) seed = std::strtoull(argv[5], nullptr, 0);
    else if (const char* e = std::getenv("B200FHE_SEED")) seed = std::strtoull(e, nullptr, 0);
    else seed = std::random_device{}();
    const int bits = std::getenv("B200FHE_PRIME_BITS") ? std::atoi(std::getenv("B200FHE_PRIME_BITS")) : 50;

    std::vector<int> bit_sizes(batch, bits);
    std::vector<uint64_t> moduli(batch);
    OK(b200fhe_primes_create(dim, bit_sizes.data(), (int)batch, moduli.data()));
    b200fhe_ctx* ctx = nullptr;
    OK(b200fhe_ctx_create((int)log_dim, moduli.data(), (int)batch, 0, &ctx));

    uint64_t *h_in, *h_clean, *h_faulty;
    OK(b200fhe_host_alloc((void**)&h_in, total * 8));
    OK(b200fhe_host_alloc((void**)&h_clean, total * 8));
    OK(b200fhe_host_alloc((void**)&h_faulty, total * 8));
    std::mt19937_64 rng(seed);
    for (size_t i = 0; i < batch; i++) {
        std::uniform_int_distribution<uint64_t> dist(0, moduli[i] - 1);
        for (size_t j = 0; j < dim; j++) h_in[i * dim + j] = dist(rng);
    }
    std::memcpy(h_clean, h_in, total * 8);
    if (legacy) std::cout << "Doing NTT 2D RADIX8\n";
    OK(b200fhe_nwt_forward_host(ctx, h_clean, 1, batch, 0));

    // distinct target words, distinct bit positions inside each (ntt_test.cu:105-123)
    std::set<uint64_t> targets;
    std::uniform_int_distribution<uint64_t> pick_word(0, total - 1);
    std::uniform_int_distribution<int> pick_bit(0, 63);
    size_t n_bits = 0;
    if (legacy) {
        // n unique (word, bit) positions over the whole buffer (exp_log.txt:5 "[2D] Flipping 2 unique random bits in the data buffer...")
        std::set<uint64_t> flipped;
        std::uniform_int_distribution<uint64_t> pick_pos(0, total_bits - 1);
        while ((long long)flipped.size() < legacy_bits) flipped.insert(pick_pos(rng));
        for (uint64_t pos : flipped) h_in[pos / 64] ^= (1ULL << (pos % 64));
        std::cout << "[2D] Flipping " << flipped.size() << " unique random bits in the data buffer...\n";
        std::cout << "Doing NTT 2D RADIX8\n";
    } else if (!real_mode) {
        while ((int)targets.size() < num_syms) targets.insert(pick_word(rng));
        for (uint64_t w : targets) {
            std::set<int> pos;
            while ((int)pos.size() < num_flips) pos.insert(pick_bit(rng));
            for (int b : pos) h_in[w] ^= (1ULL << b);
            n_bits += pos.size();
        }
        std::cout << "[2D] Flipping " << n_bits << " bits across " << targets.size() << " symbols...\n";
    }
    std::memcpy(h_faulty, h_in, total * 8);
    OK(b200fhe_nwt_forward_host(ctx, h_faulty, 1, batch, 0));

    size_t hamming = 0, bad = 0;
    std::vector<size_t> where;
    for (uint64_t i = 0; i < total; i++) {
        const uint64_t d = h_clean[i] ^ h_faulty[i];
        if (d) {
            if (++bad <= 128) where.push_back(i);
            hamming += (size_t)__builtin_popcountll(d);
        }
    }
    if (hamming) {
        const double ber = (double)hamming / (double)total_bits, ser = (double)bad / (double)total;
        std::cout << "ERROR! Total bitwise Hamming distance = " << hamming << " (bit error rate = " << ber << ")\n";
        std::cout << "       Affected symbols = " << bad << "/" << total << " (symbol error rate = " << ser << ")\n";
        std::fprintf(stderr, "[FAULT DETECTED] Bit error: %zu/%zu = %.6f, Symbol error: %zu/%zu = %.6f\n", hamming,
                     (size_t)total_bits, ber, bad, (size_t)total, ser);
    } else {
        std::cout << "ALL CORRECT\n";
    }
    if (bad > 0 && bad <= 128) {
        std::cout << "[Debug] Symbol mismatch indices:\n";
        for (size_t i : where) std::cout << "  - Index " << i << "\n";
    }
    b200fhe_host_free(h_in); b200fhe_host_free(h_clean); b200fhe_host_free(h_faulty);
    b200fhe_ctx_destroy(ctx);
    return 0;
}
