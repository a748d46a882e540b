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


def _is_prime(n):
    if n < 2:
        return False
    for p in (2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37):
        if n % p == 0:
            return n == p
    d, s = n - 1, 0
    while d % 2 == 0:
        d //= 2; s += 1
    for a in (2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37):
        x = pow(a, d, n)
        if x in (1, n - 1):
            continue
        for _ in range(s - 1):
            x = x * x % n
            if x == n - 1:
                break
        else:
            return False
    return True


def _primes_far_from_a_power_of_two(N, start, count):
    out, k = [], start // (2 * N)
    while len(out) < count:
        if _is_prime(k * 2 * N + 1):
            out.append(k * 2 * N + 1)
        k -= 1
    return out


@pytest.mark.parametrize("log_n", [12, 14, 16, 17])
def test_general_moduli_take_the_non_pseudo_mersenne_kernels(pkg, O, log_n):
    """CoeffModulus::Create only yields primes just below 2^bits, which run the NEAR kernels
    (q = 2^bits - delta).  Arbitrary NTT primes (anywhere below 2^60, e.g. the special primes of a
    key-switching basis) must run the general kernels: alone, and mixed with NEAR primes in one call."""
    N = 1 << log_n
    far = _primes_far_from_a_power_of_two(N, int(0.71 * 2 ** 60), 2) + _primes_far_from_a_power_of_two(N, int(0.6 * 2 ** 47), 1)
    near = pkg.primes_create(N, [60])
    for moduli in (far, far[:2] + near):
        ctx = pkg.Context(log_n, moduli)
        a = O.uniform_limbs(moduli, N, seed=77 + log_n)
        tabs = [O.Tables(log_n, q) for q in moduli]
        got = ctx.nwt_forward_host(a.copy())
        assert np.array_equal(got, O.ntt_forward_batch(a, tabs))
        assert np.array_equal(ctx.nwt_backward_host(got.copy()), a)
        wild = np.random.default_rng(log_n).integers(0, 2 ** 64, size=a.shape, dtype=np.uint64)
        assert np.array_equal(ctx.nwt_forward_host(wild.copy()), O.ntt_forward_batch(wild, tabs))
        _, verdict, sums = ctx.nwt_forward_protected_host(a.copy())
        assert not verdict.any() and (sums[:, 0] == sums[:, 1]).all()


def test_many_polys_share_twiddle_tiles(pkg, O):
    """The grid runs the polynomials of a batch back to back per modulus (L2 reuse of the twiddle
    tile): 5 polys x 3 limbs at N=2^16, every limb against the oracle."""
    log_n, L, n_poly = 16, 3, 5
    ctx, moduli = _ctx(pkg, log_n, 60, L)
    N = 1 << log_n
    tabs = [O.Tables(log_n, q) for q in moduli]
    a = np.stack([O.uniform_limbs(moduli, N, seed=300 + i) for i in range(n_poly)])
    want = np.stack([O.ntt_forward_batch(a[i], tabs) for i in range(n_poly)])
    got = ctx.nwt_forward_host(a.copy())
    assert np.array_equal(got, want)
    assert np.array_equal(ctx.nwt_backward_host(got.copy()), a)
    _, verdict, _ = ctx.nwt_forward_protected_host(a.copy())
    assert not verdict.any()
