// pmult_fault_test -- RNS ciphertext x plaintext multiply-accumulate under injected bit flips, on libb200fhe.so.
//
// A REDUCED experiment, not the reference's dotprod_test: that program (reliability_test/dotprod_test.cu, a whole BGV pipeline)
// is built UNMODIFIED against shim/ into oracle/_ref/dotprod_test_shim (harness/Makefile `ref`, tests/test_bgv_shim_gpu.py).
// This binary (named dotprod_test until round 2; renamed so the reference's name only ever yields the reference's report) keeps
// that program's CLI (:190-196 `<bits_per_symbol> <num_symbols>`) and injection lines (:38-61, stderr
// "Injected bitflip @ idx=..., bit=...") and its element-wise symbol / Hamming report (:118-140), but exercises only the part on
// the streaming hot path: the NTT-domain product-accumulate
//     out[p][l][i] = sum_k ct_k[p][l][i] * pt_k[l][i] mod q_l        (config C3,
//     lattigo-error-demo/seal_pmult.cpp:52, profile_framewk/src/matmul_ckks.cpp:90-107)
// on ciphertexts resident in device memory, flips applied by the device-side single-bit kernel as dotprod_test.cu:55 does,
// clean result vs faulty result.  Flips are drawn over all 2 L N words of the ciphertext; the error denominator is the output size.
//   pmult_fault_test <bits_per_symbol> <num_symbols> [log_n=14] [L=6] [terms=64] [seed]
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <random>
#include <vector>
#include "b200fhe.h"

static void die(const char* what) {
    std::fprintf(stderr, "ERROR: %s: %s\n", what, b200fhe_last_error());
    std::exit(2);
}
#define OK(call) do { if ((call) != 0) die(#call); } while (0)
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { std::fprintf(stderr, "CUDA: %s\n", cudaGetErrorString(e_)); std::exit(2); } } while (0)

int main(int argc, char** argv) {
    if (argc < 3) {
        std::cerr << "Usage: " << argv[0] << " <bits_per_symbol> <num_symbols>\n";
        return 1;
    }
    const int bits_per_symbol = std::atoi(argv[1]), num_symbols = std::atoi(argv[2]);
    const int log_n = argc > 3 ? std::atoi(argv[3]) : 14;      // dotprod_test.cu:199 N = 16384
    const int L = argc > 4 ? std::atoi(argv[4]) : 6;           // :200 six 50-bit primes
    const int T = argc > 5 ? std::atoi(argv[5]) : 64;
    const uint64_t seed = argc > 6 ? std::strtoull(argv[6], nullptr, 0) : std::random_device{}();
    if (bits_per_symbol < 0 || bits_per_symbol > 64 || num_symbols < 0 || log_n < 1 || L < 1 || T < 1) {
        std::cerr << "Error: bad arguments\n";
        return 1;
    }
    const size_t N = (size_t)1 << log_n, P = 2, ptw = (size_t)L * N, ctw = P * ptw;
    std::vector<int> bits(L, 50);
    std::vector<uint64_t> q(L);
    OK(b200fhe_primes_create(N, bits.data(), L, q.data()));
    b200fhe_ctx* ctx = nullptr;
    OK(b200fhe_ctx_create(log_n, q.data(), L, 0, &ctx));
    std::cout << "/ scheme: RNS ct x pt dot product | poly_modulus_degree: " << N << " | coeff_modulus: " << 50 * L
              << " (" << L << " x 50) bits | terms: " << T << "\n\n";

    // synthetic NTT-domain ciphertexts and plaintexts, uniform residues
    std::mt19937_64 rng(seed);
    std::vector<uint64_t> h_ct((size_t)T * ctw), h_pt((size_t)T * ptw);
    for (int k = 0; k < T; k++)
        for (size_t p = 0; p < P; p++)
            for (int l = 0; l < L; l++) {
                std::uniform_int_distribution<uint64_t> d(0, q[l] - 1);
                for (size_t i = 0; i < N; i++) h_ct[(size_t)k * ctw + (p * L + l) * N + i] = d(rng);
            }
    for (int k = 0; k < T; k++)
        for (int l = 0; l < L; l++) {
            std::uniform_int_distribution<uint64_t> d(0, q[l] - 1);
            for (size_t i = 0; i < N; i++) h_pt[(size_t)k * ptw + (size_t)l * N + i] = d(rng);
        }
    uint64_t *d_ct, *d_pt, *d_out;
    CU(cudaMalloc(&d_ct, h_ct.size() * 8));
    CU(cudaMalloc(&d_pt, h_pt.size() * 8));
    CU(cudaMalloc(&d_out, ctw * 8));
    CU(cudaMemcpy(d_ct, h_ct.data(), h_ct.size() * 8, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_pt, h_pt.data(), h_pt.size() * 8, cudaMemcpyHostToDevice));
    std::vector<const uint64_t*> pc(T), pp(T);
    for (int k = 0; k < T; k++) { pc[k] = d_ct + (size_t)k * ctw; pp[k] = d_pt + (size_t)k * ptw; }

    std::vector<uint64_t> clean(ctw), faulty(ctw);
    OK(b200fhe_pmult_accum(ctx, pc.data(), pp.data(), T, d_out, P, L, 0, 0, nullptr));
    CU(cudaMemcpy(clean.data(), d_out, ctw * 8, cudaMemcpyDeviceToHost));

    // the flips land in ciphertext 0 (x_cipher of dotprod_test.cu:109), device side
    std::uniform_int_distribution<size_t> pick(0, ctw - 1);
    std::uniform_int_distribution<unsigned> pick_bit(0, 63);
    for (int s = 0; s < num_symbols; s++) {
        const size_t idx = pick(rng);
        for (int b = 0; b < bits_per_symbol; b++) {
            const unsigned bit = pick_bit(rng);
            OK(b200fhe_flip_bit(d_ct, idx, bit, nullptr));
            CU(cudaDeviceSynchronize());
            std::cerr << "Injected bitflip @ idx=" << idx << ", bit=" << bit << "\n";
        }
    }
    OK(b200fhe_pmult_accum(ctx, pc.data(), pp.data(), T, d_out, P, L, 0, 0, nullptr));
    CU(cudaMemcpy(faulty.data(), d_out, ctw * 8, cudaMemcpyDeviceToHost));

    size_t symbol_errors = 0, bit_errors = 0;
    for (size_t i = 0; i < ctw; i++)
        if (clean[i] != faulty[i]) {
            symbol_errors++;
            bit_errors += (size_t)__builtin_popcountll(clean[i] ^ faulty[i]);
        }
    std::cout << "Elementwise symbol errors: " << symbol_errors << " / " << ctw << std::endl;
    std::cout << "Elementwise Hamming distance (bit errors): " << bit_errors << std::endl;
    if (symbol_errors == 0) std::cout << "✔ Dot product matches clean result." << std::endl;
    else std::cout << "✖ MISMATCH detected!" << std::endl;
    cudaFree(d_ct); cudaFree(d_pt); cudaFree(d_out);
    b200fhe_ctx_destroy(ctx);
    return 0;
}
