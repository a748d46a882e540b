"""GPU parity: CUDA NTT/INTT through the C ABI vs the CPU oracle, bit-exact.

Mirrors what reliability_test/ntt_test.cu does (forward NTT of a [batch][N]
buffer, compare words) with the oracle standing in for the clean reference.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ctx(pkg, log_n, bits, L):
    moduli = pkg.primes_create(1 << log_n, [bits] * L)
    return pkg.Context(log_n, moduli), moduli


@pytest.mark.parametrize("log_n,bits,L", [
    (4, 30, 2), (6, 20, 3), (8, 60, 2), (10, 50, 2), (11, 60, 3),
    (12, 60, 1), (12, 50, 32), (13, 60, 3), (14, 50, 6), (15, 60, 2), (16, 60, 4), (17, 50, 1), (17, 60, 2),
])
def test_forward_inverse_vs_oracle(pkg, O, log_n, bits, L):
    ctx, moduli = _ctx(pkg, log_n, bits, L)
    N = 1 << log_n
    a = O.uniform_limbs(moduli, N, seed=1000 + log_n)
    tabs = [O.Tables(log_n, q) for q in moduli]
    want = O.ntt_forward_batch(a, tabs)
    got = ctx.nwt_forward_host(a.copy())
    assert np.array_equal(got, want)
    back = ctx.nwt_backward_host(got.copy())
    assert np.array_equal(back, a)
    # inverse on its own against the oracle's inverse of random (bit-reversed domain) data
    b = O.uniform_limbs(moduli, N, seed=2000 + log_n)
    assert np.array_equal(ctx.nwt_backward_host(b.copy()), O.ntt_inverse_batch(b, tabs))


@pytest.mark.parametrize("log_n", [8, 12, 16])
def test_edge_inputs(pkg, O, log_n):
    """0, q-1 everywhere, and unit impulses (the reference tests no edge cases;
    these are the NTT's own: all-zero, max residue, delta at first/last index)."""
    ctx, moduli = _ctx(pkg, log_n, 60, 2)
    N = 1 << log_n
    tabs = [O.Tables(log_n, q) for q in moduli]
    cases = [np.zeros((2, N), dtype=np.uint64),
             np.stack([np.full(N, q - 1, dtype=np.uint64) for q in moduli])]
    for pos in (0, 1, N // 2, N - 1):
        d = np.zeros((2, N), dtype=np.uint64); d[:, pos] = 1
        cases.append(d)
    for a in cases:
        want = O.ntt_forward_batch(a, tabs)
        assert np.array_equal(ctx.nwt_forward_host(a.copy()), want)
        assert np.array_equal(ctx.nwt_backward_host(want.copy()), a)


@pytest.mark.parametrize("log_n,bits", [(12, 50), (16, 60), (16, 50)])
def test_out_of_range_words_are_taken_mod_q(pkg, O, log_n, bits):
    """ntt_test.cu:105-135 flips arbitrary bits 0..63, so words can exceed q;
    documented semantics: every loaded word is taken mod q."""
    ctx, moduli = _ctx(pkg, log_n, bits, 2)
    N = 1 << log_n
    rng = np.random.default_rng(5)
    a = rng.integers(0, 2**64, size=(2, N), dtype=np.uint64)
    tabs = [O.Tables(log_n, q) for q in moduli]
    assert np.array_equal(ctx.nwt_forward_host(a.copy()), O.ntt_forward_batch(a, tabs))
    assert np.array_equal(ctx.nwt_backward_host(a.copy()), O.ntt_inverse_batch(a, tabs))


def test_multi_poly_and_start_mod(pkg, O):
    """[n_poly][L][N] batches and Phantom's start_modulus_idx."""
    log_n, Lctx = 12, 6
    ctx, moduli = _ctx(pkg, log_n, 60, Lctx)
    N = 1 << log_n
    tabs = [O.Tables(log_n, q) for q in moduli]
    start, L, n_poly = 2, 3, 4
    sub = moduli[start:start + L]
    a = np.stack([O.uniform_limbs(sub, N, seed=i) for i in range(n_poly)])
    want = np.stack([O.ntt_forward_batch(a[i], tabs[start:start + L]) for i in range(n_poly)])
    got = ctx.nwt_forward_host(a.copy(), L=L, start_mod=start)
    assert np.array_equal(got, want)
    assert np.array_equal(ctx.nwt_backward_host(got.copy(), L=L, start_mod=start), a)


def test_full_size_properties(pkg, O):
    """BASELINE config 2 shape (N=2^16, L=36): round trip + linearity + the
    negacyclic convolution theorem against a sparse schoolbook product."""
    log_n, L = 16, 36
    ctx, moduli = _ctx(pkg, log_n, 60, L)
    N = 1 << log_n
    a = O.uniform_limbs(moduli, N, seed=11)
    b = O.uniform_limbs(moduli, N, seed=12)
    A = ctx.nwt_forward_host(a.copy()); B = ctx.nwt_forward_host(b.copy())
    assert np.array_equal(ctx.nwt_backward_host(A.copy()), a)
    qs = np.array(moduli, dtype=object)[:, None]
    s = ((a.astype(object) + b.astype(object)) % qs).astype(np.uint64)
    S = ctx.nwt_forward_host(s.copy())
    assert np.array_equal(S, ((A.astype(object) + B.astype(object)) % qs).astype(np.uint64))
    # x^k * a(x) mod x^N+1 via the transform: multiply by NTT(x^k) pointwise
    k = 12345
    mono = np.zeros((L, N), dtype=np.uint64); mono[:, k] = 1
    M = ctx.nwt_forward_host(mono.copy())
    prod = ((A.astype(object) * M.astype(object)) % qs).astype(np.uint64)
    c = ctx.nwt_backward_host(prod)
    want = np.empty_like(a)
    want[:, k:] = a[:, :N - k]
    want[:, :k] = ((qs - a[:, N - k:].astype(object)) % qs).astype(np.uint64)
    assert np.array_equal(c, want)
    # spot-check against the oracle on 3 limbs
    for l in (0, 17, 35):
        t = O.Tables(log_n, moduli[l])
        assert np.array_equal(A[l], O.ntt_forward(a[l], t))


def test_error_paths(pkg):
    import ctypes as C
    with pytest.raises(pkg.B200FHEError):
        pkg.Context(12, [12345])            # not prime / not 1 mod 2N
    with pytest.raises(pkg.B200FHEError):
        pkg.Context(12, [(1 << 61) - 1])    # too large
    ctx, _ = _ctx(pkg, 12, 60, 2)
    a = np.zeros((3, 1 << 12), dtype=np.uint64)
    with pytest.raises(pkg.B200FHEError):
        ctx.nwt_forward_host(a, L=3, start_mod=0)   # exceeds context
    # empty batch is a no-op
    assert pkg.lib().b200fhe_nwt_forward(ctx._h, None, 0, 2, 0, None) == 0
