// Source-compatible stand-in for the subset of phantom-fhe's "phantom.h" that
// reliability_test/ntt_test.cu and ntt_real_test.cu use (reference lines cited per item).
// Header-only; every call forwards to the C ABI of libb200fhe.so (include/b200fhe.h), so the
// reference harness compiles UNMODIFIED with `-I shim -I include` and links `-lb200fhe`
// instead of `-lPhantom`.  Names, argument order and ownership follow the
// call sites; nothing here is copied from phantom-fhe (its sources are not in the reference tree).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include <vector>
#include "b200fhe.h"

namespace phantom {
namespace detail {
inline void check(int rc, const char* what) {
    if (rc != 0) throw std::logic_error(std::string(what) + ": " + b200fhe_last_error());
}
}  // namespace detail

namespace util {
// ntt_test.cu:40-41 -- `cuda_stream_wrapper w; const auto& s = w.get_stream();`
class cuda_stream_wrapper {
public:
    cuda_stream_wrapper() {
        if (cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking) != cudaSuccess)
            throw std::runtime_error("cuda_stream_wrapper: no usable CUDA device (this library has no CPU fallback)");
    }
    ~cuda_stream_wrapper() { if (stream_) cudaStreamDestroy(stream_); }
    cuda_stream_wrapper(const cuda_stream_wrapper&) = delete;
    cuda_stream_wrapper& operator=(const cuda_stream_wrapper&) = delete;
    const cudaStream_t& get_stream() const { return stream_; }
private:
    cudaStream_t stream_ = nullptr;
};

// ntt_test.cu:47,88 -- `auto p = make_cuda_auto_ptr<T>(n, s); p.get()`.  The harness writes
// DModulus entries through get() on the HOST (ntt_test.cu:49) and hands the same pointer to
// kernels, so the allocation is managed memory: valid on both sides.  Freed on scope exit.
template <typename T>
class cuda_auto_ptr {
public:
    cuda_auto_ptr() = default;
    cuda_auto_ptr(size_t n, const cudaStream_t& s) : n_(n), s_(s) {
        if (n && cudaMallocManaged(reinterpret_cast<void**>(&p_), n * sizeof(T)) != cudaSuccess)
            throw std::bad_alloc();
    }
    cuda_auto_ptr(cuda_auto_ptr&& o) noexcept : p_(o.p_), n_(o.n_), s_(o.s_) { o.p_ = nullptr; }
    cuda_auto_ptr& operator=(cuda_auto_ptr&& o) noexcept {
        if (this != &o) { reset(); p_ = o.p_; n_ = o.n_; s_ = o.s_; o.p_ = nullptr; }
        return *this;
    }
    cuda_auto_ptr(const cuda_auto_ptr&) = delete;
    cuda_auto_ptr& operator=(const cuda_auto_ptr&) = delete;
    ~cuda_auto_ptr() { reset(); }
    T* get() const { return p_; }
    size_t get_n() const { return n_; }
    void reset() {
        if (p_) { cudaStreamSynchronize(s_); cudaFree(p_); p_ = nullptr; }
    }
private:
    T* p_ = nullptr;
    size_t n_ = 0;
    cudaStream_t s_ = nullptr;
};
template <typename T>
inline cuda_auto_ptr<T> make_cuda_auto_ptr(size_t n, const cudaStream_t& s) { return cuda_auto_ptr<T>(n, s); }
}  // namespace util

namespace arith {
// ntt_test.cu:44,50-52 -- value(), const_ratio()[0..1] (floor(2^128/q), two words).
class Modulus {
public:
    Modulus() = default;
    explicit Modulus(uint64_t v) { set_value(v); }
    void set_value(uint64_t v) {
        value_ = v;
        ratio_[0] = ratio_[1] = ratio_[2] = 0;
        if (v > 1) {
            detail::check(b200fhe_const_ratio(v, ratio_), "Modulus::set_value");
            unsigned __int128 two64 = (unsigned __int128)1 << 64;   // remainder word, as SEAL keeps it
            unsigned __int128 rem = ((two64 % v) * (two64 % v)) % v;
            ratio_[2] = (uint64_t)rem;
        }
    }
    uint64_t value() const { return value_; }
    const uint64_t* const_ratio() const { return ratio_; }
    int bit_count() const { int b = 0; for (uint64_t v = value_; v; v >>= 1) b++; return b; }
    bool is_zero() const { return value_ == 0; }
private:
    uint64_t value_ = 0;
    uint64_t ratio_[3] = {0, 0, 0};
};

// ntt_test.cu:44, dotprod_test.cu:200 -- CoeffModulus::Create(N, bit_sizes).
struct CoeffModulus {
    static std::vector<Modulus> Create(size_t poly_modulus_degree, const std::vector<int>& bit_sizes) {
        std::vector<uint64_t> q(bit_sizes.size());
        if (!bit_sizes.empty() &&
            b200fhe_primes_create(poly_modulus_degree, bit_sizes.data(), (int)bit_sizes.size(), q.data()) != 0)
            throw std::logic_error(std::string("CoeffModulus::Create: ") + b200fhe_last_error());
        std::vector<Modulus> out;
        out.reserve(q.size());
        for (uint64_t v : q) out.emplace_back(v);
        return out;
    }
};

// ntt_test.cu:60,63-64 -- `NTT t(log_n, modulus); t.get_from_root_powers().data()`.
class NTT {
public:
    NTT(int log_n, const Modulus& modulus) : log_n_(log_n), modulus_(modulus) {
        const size_t n = (size_t)1 << log_n;
        rp_.resize(n); rps_.resize(n); irp_.resize(n); irps_.resize(n);
        detail::check(b200fhe_host_tables(log_n, modulus.value(), rp_.data(), rps_.data(), irp_.data(), irps_.data(),
                                          &n_inv_, &n_inv_shoup_), "NTT::NTT");
    }
    const std::vector<uint64_t>& get_from_root_powers() const { return rp_; }
    const std::vector<uint64_t>& get_from_root_powers_shoup() const { return rps_; }
    const std::vector<uint64_t>& get_from_inv_root_powers() const { return irp_; }
    const std::vector<uint64_t>& get_from_inv_root_powers_shoup() const { return irps_; }
    uint64_t inv_degree_modulo() const { return n_inv_; }
    uint64_t inv_degree_modulo_shoup() const { return n_inv_shoup_; }
    int coeff_count_power() const { return log_n_; }
    const Modulus& modulus() const { return modulus_; }
private:
    int log_n_;
    Modulus modulus_;
    std::vector<uint64_t> rp_, rps_, irp_, irps_;
    uint64_t n_inv_ = 0, n_inv_shoup_ = 0;
};
}  // namespace arith
}  // namespace phantom

// ---- global-namespace device-side types of Phantom used by the harness ----------------------

// ntt_test.cu:47-54 -- `modulus.get()[i].set(value, const_ratio[0], const_ratio[1])`, written from
// the host into make_cuda_auto_ptr memory.  Plain data; the kernels of libb200fhe.so keep their own
// per-modulus constants, this struct only carries the modulus to DNTTTable::set.
struct DModulus {
    uint64_t value_ = 0;
    uint64_t const_ratio_[2] = {0, 0};
    void set(uint64_t value, uint64_t ratio0, uint64_t ratio1) {
        value_ = value;
        const_ratio_[0] = ratio0;
        const_ratio_[1] = ratio1;
    }
    uint64_t value() const { return value_; }
    const uint64_t* const_ratio() const { return const_ratio_; }
};

// ntt_test.cu:57-68 -- `DNTTTable t; t.init(N, L, s); t.set(&modulus[i], tw, tw_shoup, itw,
// itw_shoup, n_inv, n_inv_shoup, i, s);`.  Owns a b200fhe context; host twiddles are copied by set
// (the host NTT object dies at the end of each loop iteration, ntt_test.cu:60-69).
class DNTTTable {
public:
    DNTTTable() = default;
    DNTTTable(const DNTTTable&) = delete;
    DNTTTable& operator=(const DNTTTable&) = delete;
    ~DNTTTable() { if (ctx_) b200fhe_ctx_destroy(ctx_); }

    void init(size_t n, size_t size, const cudaStream_t& /*stream*/) {
        if (ctx_) { b200fhe_ctx_destroy(ctx_); ctx_ = nullptr; }
        int log_n = 0;
        while (((size_t)1 << log_n) < n) log_n++;
        if (((size_t)1 << log_n) != n) throw std::invalid_argument("DNTTTable::init: n must be a power of two");
        int dev = 0;
        cudaGetDevice(&dev);
        phantom::detail::check(b200fhe_ctx_create_empty(log_n, (int)size, dev, &ctx_), "DNTTTable::init");
        n_ = n;
        size_ = size;
    }
    void set(const DModulus* modulus, const uint64_t* twiddle, const uint64_t* twiddle_shoup, const uint64_t* itwiddle,
             const uint64_t* itwiddle_shoup, uint64_t n_inv_mod_q, uint64_t n_inv_mod_q_shoup, size_t index,
             const cudaStream_t& stream) {
        if (!ctx_ || !modulus) throw std::invalid_argument("DNTTTable::set: table not initialised");
        phantom::detail::check(b200fhe_ctx_set_tables(ctx_, (int)index, modulus->value(), twiddle, twiddle_shoup, itwiddle,
                                                      itwiddle_shoup, n_inv_mod_q, n_inv_mod_q_shoup, (void*)stream),
                               "DNTTTable::set");
    }
    size_t n() const { return n_; }
    size_t size() const { return size_; }
    b200fhe_ctx* ctx() const { return ctx_; }
private:
    b200fhe_ctx* ctx_ = nullptr;
    size_t n_ = 0, size_ = 0;
};

// ntt_test.cu:95,144; ntt_real_test.cu:89,126.  In-place forward negacyclic NTT of
// coeff_modulus_size limbs, limb i under modulus start_modulus_idx + i.  Errors throw, as
// upstream Phantom's argument checks do; the harness never checks CUDA status itself.
inline void nwt_2d_radix8_forward_inplace(uint64_t* inout, const DNTTTable& ntt_tables, size_t coeff_modulus_size,
                                          size_t start_modulus_idx, const cudaStream_t& stream) {
    phantom::detail::check(b200fhe_nwt_forward(ntt_tables.ctx(), inout, 1, coeff_modulus_size, start_modulus_idx, (void*)stream),
                           "nwt_2d_radix8_forward_inplace");
}
// Phantom's matching inverse (reached inside multiply/relinearize/decrypt, dotprod_test.cu:113-119).
inline void nwt_2d_radix8_backward_inplace(uint64_t* inout, const DNTTTable& ntt_tables, size_t coeff_modulus_size,
                                           size_t start_modulus_idx, const cudaStream_t& stream) {
    phantom::detail::check(b200fhe_nwt_backward(ntt_tables.ctx(), inout, 1, coeff_modulus_size, start_modulus_idx, (void*)stream),
                           "nwt_2d_radix8_backward_inplace");
}

// ---- BGV class surface used by dotprod_test.cu / dotprod_real_test.cu / naive_gemm_test.cu ----
#include "phantom_bgv.h"
