// Source-compatible stand-in for the BGV class surface of phantom-fhe that reliability_test/dotprod_test.cu,
// dotprod_real_test.cu and naive_gemm_test.cu use (call sites cited per item; symbol list SURVEY 8(b)).  Header-only;
// every method forwards to the b200fhe_bgv_* entries of libb200fhe.so (include/b200fhe.h), where the polynomial work runs
// on the GPU.  Included by shim/phantom.h; names, argument order and value semantics follow the call sites, nothing here
// is copied from phantom-fhe (its sources are not in the reference tree).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdlib>
#include <memory>
#include <random>
#include <stdexcept>
#include <string>
#include <vector>
#include "b200fhe.h"

namespace phantom {

// dotprod_test.cu:26 -- `EncryptionParameters parms(scheme_type::bgv);` (a GLOBAL: the constructor must not touch CUDA)
enum class scheme_type : uint8_t { none = 0, bgv = 1, bfv = 2, ckks = 3 };

namespace arith {
// dotprod_test.cu:206 -- PlainModulus::Batching(N, 20): the largest 20-bit prime = 1 mod 2N
struct PlainModulus {
    static Modulus Batching(size_t poly_modulus_degree, int bit_size) {
        return CoeffModulus::Create(poly_modulus_degree, std::vector<int>{bit_size})[0];
    }
};
}  // namespace arith

// dotprod_test.cu:199-206
class EncryptionParameters {
public:
    explicit EncryptionParameters(scheme_type scheme = scheme_type::bgv) : scheme_(scheme) {}
    void set_poly_modulus_degree(size_t n) { n_ = n; }
    void set_coeff_modulus(const std::vector<arith::Modulus>& m) { coeff_ = m; }
    void set_special_modulus_size(size_t k) { special_ = k; }
    void set_plain_modulus(const arith::Modulus& t) { plain_ = t; }
    void set_plain_modulus(uint64_t t) { plain_ = arith::Modulus(t); }
    scheme_type scheme() const { return scheme_; }
    size_t poly_modulus_degree() const { return n_; }
    const std::vector<arith::Modulus>& coeff_modulus() const { return coeff_; }
    size_t special_modulus_size() const { return special_; }
    const arith::Modulus& plain_modulus() const { return plain_; }
private:
    scheme_type scheme_;
    size_t n_ = 0, special_ = 1;
    std::vector<arith::Modulus> coeff_;
    arith::Modulus plain_;
};
}  // namespace phantom

// dotprod_test.cu:209 -- `PhantomContext context(parms);`
class PhantomContext {
public:
    explicit PhantomContext(const phantom::EncryptionParameters& parms) : parms_(parms) {
        if (parms.scheme() != phantom::scheme_type::bgv) throw std::invalid_argument("PhantomContext: this library implements the BGV path");
        int log_n = 0;
        while (((size_t)1 << log_n) < parms.poly_modulus_degree()) log_n++;
        if (((size_t)1 << log_n) != parms.poly_modulus_degree()) throw std::invalid_argument("PhantomContext: poly_modulus_degree must be a power of two");
        std::vector<uint64_t> q;
        for (const auto& m : parms.coeff_modulus()) q.push_back(m.value());
        uint64_t seed;
        if (const char* s = std::getenv("B200FHE_SEED")) seed = std::strtoull(s, nullptr, 10);
        else { std::random_device rd; seed = ((uint64_t)rd() << 32) ^ rd(); }
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) throw std::runtime_error("PhantomContext: no usable CUDA device (this library has no CPU fallback)");
        b200fhe_bgv* h = nullptr;
        phantom::detail::check(b200fhe_bgv_create(log_n, q.data(), (int)q.size(), (int)parms.special_modulus_size(), parms.plain_modulus().value(),
                                                  seed, dev, &h), "PhantomContext");
        h_.reset(h, [](b200fhe_bgv* p) { b200fhe_bgv_destroy(p); });
        n_q_ = q.size() - parms.special_modulus_size();
    }
    const phantom::EncryptionParameters& parms() const { return parms_; }
    b200fhe_bgv* bgv() const { return h_.get(); }
    size_t poly_modulus_degree() const { return parms_.poly_modulus_degree(); }
    size_t first_coeff_modulus_size() const { return n_q_; }
private:
    phantom::EncryptionParameters parms_;
    std::shared_ptr<b200fhe_bgv> h_;
    size_t n_q_ = 0;
};

// dotprod_test.cu:100-101 -- the encoder's output / the decryptor's result: N coefficients mod t on the host
class PhantomPlaintext {
public:
    std::vector<uint64_t>& coeffs() { return c_; }
    const std::vector<uint64_t>& coeffs() const { return c_; }
    size_t coeff_count() const { return c_.size(); }
    uint64_t* data() { return c_.data(); }
    const uint64_t* data() const { return c_.data(); }
private:
    std::vector<uint64_t> c_;
};

// dotprod_test.cu:42-46,104,144 -- device-resident ciphertext [size][coeff_modulus_size][poly_modulus_degree], NTT form;
// data() is the raw device pointer the harness flips bits in; copyable (`PhantomCiphertext rotated = xy_cipher;`).
class PhantomCiphertext {
public:
    PhantomCiphertext() = default;
    PhantomCiphertext(const PhantomCiphertext& o) { *this = o; }
    PhantomCiphertext(PhantomCiphertext&& o) noexcept { swap(o); }
    PhantomCiphertext& operator=(const PhantomCiphertext& o) {
        if (this != &o) {
            resize(o.size_, o.limbs_, o.n_);
            correction_ = o.correction_;
            if (o.words() && cudaMemcpy(d_, o.d_, o.words() * sizeof(uint64_t), cudaMemcpyDeviceToDevice) != cudaSuccess)
                throw std::runtime_error("PhantomCiphertext: device copy failed");
        }
        return *this;
    }
    PhantomCiphertext& operator=(PhantomCiphertext&& o) noexcept { swap(o); return *this; }
    ~PhantomCiphertext() { if (d_) cudaFree(d_); }
    void resize(size_t size, size_t limbs, size_t n) {
        if (size * limbs * n != words()) {
            if (d_) { cudaFree(d_); d_ = nullptr; }
            if (size * limbs * n && cudaMalloc(reinterpret_cast<void**>(&d_), size * limbs * n * sizeof(uint64_t)) != cudaSuccess) throw std::bad_alloc();
        }
        size_ = size; limbs_ = limbs; n_ = n;
    }
    size_t size() const { return size_; }
    size_t coeff_modulus_size() const { return limbs_; }
    size_t poly_modulus_degree() const { return n_; }
    uint64_t* data() const { return d_; }
    uint64_t correction_factor() const { return correction_; }
    void set_correction_factor(uint64_t c) { correction_ = c; }
    void swap(PhantomCiphertext& o) noexcept {
        std::swap(d_, o.d_); std::swap(size_, o.size_); std::swap(limbs_, o.limbs_); std::swap(n_, o.n_); std::swap(correction_, o.correction_);
    }
private:
    size_t words() const { return size_ * limbs_ * n_; }
    uint64_t* d_ = nullptr;
    size_t size_ = 0, limbs_ = 0, n_ = 0;
    uint64_t correction_ = 1;
};

// dotprod_test.cu:75,143 -- handles: the switching keys themselves are generated on first use, per level, inside the
// library object that holds the secret key (include/b200fhe.h, BGV section); the handle keeps that object alive.
class PhantomRelinKey {
public:
    PhantomRelinKey() = default;
    explicit PhantomRelinKey(const PhantomContext& c) : bgv_(c.bgv()) {}
    b200fhe_bgv* bgv() const { return bgv_; }
private:
    b200fhe_bgv* bgv_ = nullptr;
};
class PhantomGaloisKey {
public:
    PhantomGaloisKey() = default;
    explicit PhantomGaloisKey(const PhantomContext& c) : bgv_(c.bgv()) {}
    b200fhe_bgv* bgv() const { return bgv_; }
private:
    b200fhe_bgv* bgv_ = nullptr;
};

// dotprod_test.cu:74,105-106
class PhantomPublicKey {
public:
    PhantomPublicKey() = default;
    explicit PhantomPublicKey(const PhantomContext& c) : bgv_(c.bgv()) {}
    void encrypt_asymmetric(const PhantomContext& context, const PhantomPlaintext& plain, PhantomCiphertext& cipher) const {
        if (!bgv_ || bgv_ != context.bgv()) throw std::invalid_argument("encrypt_asymmetric: key does not belong to this context");
        if (plain.coeff_count() != context.poly_modulus_degree()) throw std::invalid_argument("encrypt_asymmetric: plaintext size");
        cipher.resize(2, context.first_coeff_modulus_size(), context.poly_modulus_degree());
        cipher.set_correction_factor(1);
        phantom::detail::check(b200fhe_bgv_encrypt(bgv_, plain.data(), cipher.data()), "encrypt_asymmetric");
    }
    PhantomCiphertext encrypt_asymmetric(const PhantomContext& context, const PhantomPlaintext& plain) const {
        PhantomCiphertext c;
        encrypt_asymmetric(context, plain, c);
        return c;
    }
private:
    b200fhe_bgv* bgv_ = nullptr;
};

// dotprod_test.cu:73-75,118,143,152
class PhantomSecretKey {
public:
    explicit PhantomSecretKey(const PhantomContext& context) : bgv_(context.bgv()) { gen_secretkey(context); }
    void gen_secretkey(const PhantomContext& context) {
        phantom::detail::check(b200fhe_bgv_gen_secretkey(context.bgv()), "gen_secretkey");
    }
    PhantomPublicKey gen_publickey(const PhantomContext& context) const {
        phantom::detail::check(b200fhe_bgv_gen_publickey(context.bgv()), "gen_publickey");
        return PhantomPublicKey(context);
    }
    PhantomRelinKey gen_relinkey(const PhantomContext& context) const { return PhantomRelinKey(context); }
    PhantomGaloisKey create_galois_keys(const PhantomContext& context) const { return PhantomGaloisKey(context); }
    void decrypt(const PhantomContext& context, const PhantomCiphertext& cipher, PhantomPlaintext& plain) const {
        plain.coeffs().assign(context.poly_modulus_degree(), 0);
        phantom::detail::check(b200fhe_bgv_decrypt(context.bgv(), cipher.data(), (int)cipher.size(), (int)cipher.coeff_modulus_size(),
                                                   cipher.correction_factor(), plain.data()), "decrypt");
    }
    PhantomPlaintext decrypt(const PhantomContext& context, const PhantomCiphertext& cipher) const {
        PhantomPlaintext p;
        decrypt(context, cipher, p);
        return p;
    }
private:
    b200fhe_bgv* bgv_ = nullptr;
};

// dotprod_test.cu:78-80,100-101,119,153
class PhantomBatchEncoder {
public:
    explicit PhantomBatchEncoder(const PhantomContext& context) : slots_(context.poly_modulus_degree()) {}
    size_t slot_count() const { return slots_; }
    void encode(const PhantomContext& context, const std::vector<uint64_t>& values, PhantomPlaintext& plain) const {
        if (values.size() > slots_) throw std::invalid_argument("encode: more values than slots");
        plain.coeffs().assign(slots_, 0);
        phantom::detail::check(b200fhe_bgv_encode(context.bgv(), values.data(), values.size(), plain.data()), "encode");
    }
    PhantomPlaintext encode(const PhantomContext& context, const std::vector<uint64_t>& values) const {
        PhantomPlaintext p;
        encode(context, values, p);
        return p;
    }
    void decode(const PhantomContext& context, const PhantomPlaintext& plain, std::vector<uint64_t>& values) const {
        values.assign(slots_, 0);
        phantom::detail::check(b200fhe_bgv_decode(context.bgv(), plain.data(), values.data()), "decode");
    }
    std::vector<uint64_t> decode(const PhantomContext& context, const PhantomPlaintext& plain) const {
        std::vector<uint64_t> v;
        decode(context, plain, v);
        return v;
    }
private:
    size_t slots_;
};

namespace phantom {
// dotprod_test.cu:113 -- ct x ct tensor; operands at the same level
inline PhantomCiphertext multiply(const PhantomContext& context, const PhantomCiphertext& a, const PhantomCiphertext& b) {
    if (a.size() != 2 || b.size() != 2 || a.coeff_modulus_size() != b.coeff_modulus_size())
        throw std::invalid_argument("multiply: operands must be size-2 ciphertexts of the same level");
    PhantomCiphertext out;
    out.resize(3, a.coeff_modulus_size(), a.poly_modulus_degree());
    detail::check(b200fhe_bgv_multiply(context.bgv(), a.data(), b.data(), out.data(), (int)a.coeff_modulus_size()), "multiply");
    uint64_t t = 0;
    b200fhe_bgv_info(context.bgv(), nullptr, nullptr, nullptr, &t, nullptr);
    out.set_correction_factor((uint64_t)((unsigned __int128)a.correction_factor() * b.correction_factor() % t));
    return out;
}
inline void multiply_inplace(const PhantomContext& context, PhantomCiphertext& a, const PhantomCiphertext& b) {
    PhantomCiphertext out = multiply(context, a, b);
    a.swap(out);
}
// dotprod_test.cu:114
inline void relinearize_inplace(const PhantomContext& context, PhantomCiphertext& ct, const PhantomRelinKey& /*keys*/) {
    if (ct.size() == 2) return;
    if (ct.size() != 3) throw std::invalid_argument("relinearize_inplace: ciphertext size must be 3");
    detail::check(b200fhe_bgv_relinearize(context.bgv(), ct.data(), (int)ct.coeff_modulus_size()), "relinearize_inplace");
    PhantomCiphertext out;
    out.resize(2, ct.coeff_modulus_size(), ct.poly_modulus_degree());
    out.set_correction_factor(ct.correction_factor());
    if (cudaMemcpy(out.data(), ct.data(), 2 * ct.coeff_modulus_size() * ct.poly_modulus_degree() * sizeof(uint64_t), cudaMemcpyDeviceToDevice) != cudaSuccess)
        throw std::runtime_error("relinearize_inplace: device copy failed");
    ct.swap(out);
}
// dotprod_test.cu:115
inline PhantomCiphertext mod_switch_to_next(const PhantomContext& context, const PhantomCiphertext& ct) {
    PhantomCiphertext out;
    out.resize(ct.size(), ct.coeff_modulus_size() - 1, ct.poly_modulus_degree());
    uint64_t cf = ct.correction_factor();
    detail::check(b200fhe_bgv_mod_switch_to_next(context.bgv(), ct.data(), (int)ct.size(), (int)ct.coeff_modulus_size(), out.data(), &cf),
                  "mod_switch_to_next");
    out.set_correction_factor(cf);
    return out;
}
inline void mod_switch_to_next_inplace(const PhantomContext& context, PhantomCiphertext& ct) {
    PhantomCiphertext out = mod_switch_to_next(context, ct);
    ct.swap(out);
}
// dotprod_test.cu:146 -- rows rotated left by `step` slots
inline void rotate_inplace(const PhantomContext& context, PhantomCiphertext& ct, int step, const PhantomGaloisKey& /*keys*/) {
    if (ct.size() != 2) throw std::invalid_argument("rotate_inplace: relinearize first");
    detail::check(b200fhe_bgv_rotate(context.bgv(), ct.data(), (int)ct.coeff_modulus_size(), step, 0), "rotate_inplace");
}
inline void rotate_rows_inplace(const PhantomContext& context, PhantomCiphertext& ct, int step, const PhantomGaloisKey& keys) {
    rotate_inplace(context, ct, step, keys);
}
inline void rotate_columns_inplace(const PhantomContext& context, PhantomCiphertext& ct, const PhantomGaloisKey& /*keys*/) {
    detail::check(b200fhe_bgv_rotate(context.bgv(), ct.data(), (int)ct.coeff_modulus_size(), 0, 1), "rotate_columns_inplace");
}
// dotprod_test.cu:150
inline void add_inplace(const PhantomContext& context, PhantomCiphertext& a, const PhantomCiphertext& b) {
    if (a.size() != b.size() || a.coeff_modulus_size() != b.coeff_modulus_size()) throw std::invalid_argument("add_inplace: operands differ in size or level");
    if (a.correction_factor() != b.correction_factor()) throw std::invalid_argument("add_inplace: operands carry different correction factors");
    detail::check(b200fhe_bgv_add(context.bgv(), a.data(), b.data(), (int)a.size(), (int)a.coeff_modulus_size()), "add_inplace");
}
}  // namespace phantom
