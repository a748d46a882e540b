"""GPU parity: ct x pt multiply-accumulate (config C3) through the C ABI vs the oracle, bit-exact.

Shape and semantics follow lattigo-error-demo/seal_pmult.cpp:52 (multiply_plain in the NTT
domain) accumulated as in profile_framewk/src/matmul_ckks.cpp:90-107; ciphertext layout
[size][L][N] as in reliability_test/dotprod_test.cu:42-46."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rand(O, moduli, N, T, P, seed):
    ct = np.stack([np.stack([O.uniform_limbs(moduli, N, seed=seed + 31 * t + p) for p in range(P)]) for t in range(T)]) \
        if T else np.zeros((0, P, len(moduli), N), dtype=np.uint64)
    pt = np.stack([O.uniform_limbs(moduli, N, seed=seed + 7777 + t) for t in range(T)]) \
        if T else np.zeros((0, len(moduli), N), dtype=np.uint64)
    return ct, pt


@pytest.mark.parametrize("log_n,bits,L,T,P", [
    (4, [30, 20], 2, 3, 2), (10, [60, 40, 40, 60], 4, 5, 2), (12, [60] * 3, 3, 1, 1), (12, [50] * 3, 3, 64, 2),
    (13, [60, 40, 40, 60], 4, 7, 3), (12, [60] * 2, 2, 70, 2), (12, [60] * 2, 2, 0, 2), (14, [50] * 6, 6, 9, 2),
])
def test_pmult_host_vs_oracle(pkg, O, log_n, bits, L, T, P):
    N = 1 << log_n
    moduli = pkg.primes_create(N, bits)
    ctx = pkg.Context(log_n, moduli)
    ct, pt = _rand(O, moduli, N, T, P, seed=100 + log_n + T)
    got = ctx.pmult_accum_host(ct, pt)
    want = O.pmult_accum(ct, pt, moduli) if T else np.zeros((P, L, N), dtype=np.uint64)
    assert np.array_equal(got, want)


def test_pmult_out_of_range_words_are_taken_mod_q(pkg, O):
    """dotprod_test.cu:38-61 flips any of the 64 bits of a ciphertext word: words >= q must behave as word mod q."""
    log_n, N = 12, 4096
    moduli = pkg.primes_create(N, [60, 50])
    ctx = pkg.Context(log_n, moduli)
    rng = np.random.default_rng(9)
    T, P = 66, 2
    ct = rng.integers(0, 2 ** 64, size=(T, P, 2, N), dtype=np.uint64)
    pt = rng.integers(0, 2 ** 64, size=(T, 2, N), dtype=np.uint64)
    ct[0, 0, 0, :8] = np.uint64(2 ** 64 - 1)
    pt[0, 0, :8] = np.uint64(2 ** 64 - 1)
    assert np.array_equal(ctx.pmult_accum_host(ct, pt), O.pmult_accum(ct, pt, moduli))


def test_pmult_device_pointers_start_mod_and_accumulate(pkg, O):
    import torch
    log_n, N = 12, 4096
    moduli = pkg.primes_create(N, [60] * 5)
    ctx = pkg.Context(log_n, moduli)
    sub = moduli[2:5]                      # limbs 2..4 of the context
    T, P = 6, 2
    ct, pt = _rand(O, sub, N, T, P, seed=5)
    dct = [torch.from_numpy(ct[k].view(np.int64)).cuda() for k in range(T)]
    dpt = [torch.from_numpy(pt[k].view(np.int64)).cuda() for k in range(T)]
    out = torch.zeros((P, 3, N), dtype=torch.int64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    ctx.pmult_accum([t.data_ptr() for t in dct[:4]], [t.data_ptr() for t in dpt[:4]], out.data_ptr(), P, L=3, start_mod=2, stream=s)
    ctx.pmult_accum([t.data_ptr() for t in dct[4:]], [t.data_ptr() for t in dpt[4:]], out.data_ptr(), P, L=3, start_mod=2,
                    accumulate=True, stream=s)
    torch.cuda.synchronize()
    got = out.cpu().numpy().view(np.uint64)
    assert np.array_equal(got, O.pmult_accum(ct, pt, sub))


def test_pmult_linearity_at_full_size(pkg, O):
    """C3 shape (N=2^16, L=36) with 8 terms: full-size property checks (the oracle covers one limb):
    sum over all terms == sum of two half-sums (mod q), and limb 0 bit-exact vs the oracle."""
    log_n, N, L = 16, 1 << 16, 36
    moduli = pkg.primes_create(N, [60] * L)
    ctx = pkg.Context(log_n, moduli)
    T, P = 8, 2
    rng = np.random.default_rng(1)
    ct = np.empty((T, P, L, N), dtype=np.uint64); pt = np.empty((T, L, N), dtype=np.uint64)
    for l, q in enumerate(moduli):
        ct[:, :, l, :] = rng.integers(0, q, size=(T, P, N), dtype=np.uint64)
        pt[:, l, :] = rng.integers(0, q, size=(T, N), dtype=np.uint64)
    full = ctx.pmult_accum_host(ct, pt)
    a = ctx.pmult_accum_host(ct[:3], pt[:3]); b = ctx.pmult_accum_host(ct[3:], pt[3:])
    q = np.array(moduli, dtype=np.uint64)[None, :, None]
    s = a + b                                  # < 2^61, no wrap
    s = np.where(s >= q, s - q, s)
    assert np.array_equal(full, s)
    want0 = O.pmult_accum(ct[:, :, :1, :], pt[:, :1, :], moduli[:1])
    assert np.array_equal(full[:, :1, :], want0)


def test_pmult_rejects_bad_arguments(pkg):
    ctx = pkg.Context(12, pkg.primes_create(4096, [60] * 2))
    with pytest.raises(pkg.B200FHEError):
        ctx.pmult_accum([1], [1], 0, 2)                      # null output
    with pytest.raises(pkg.B200FHEError):
        ctx.pmult_accum([], [], 16, 2, L=3)                  # more limbs than the context holds
