"""Protected INVERSE NTT (b200fhe_nwt_backward_protected_host): outputs bit-exact against the oracle's inverse transform, both
checksums against an independent big-integer evaluation of the identity of rfhe_framewk/src/negaclic_ntt.py:130-149 read
backwards (sum_k w^_k a^_k over the inputs taken mod q == sum_i w_i a_i over the outputs), verdict clean; k_small (N <= 2^11),
the two-pass kernels (2^12 .. 2^17), out-of-range input words, several polynomials and limbs per call."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _dot_mod(w, x, q):
    return int(sum(int(a) * int(b) for a, b in zip(w, x)) % q)


@pytest.mark.parametrize("log_n,bits,L,n_poly", [(6, 30, 2, 3), (10, 50, 1, 2), (12, 60, 3, 2), (13, 50, 2, 1), (14, 60, 2, 1), (15, 40, 1, 2),
                                                  (16, 60, 2, 2), (17, 60, 1, 1)])
def test_protected_inverse_matches_oracle_and_identity(pkg, O, log_n, bits, L, n_poly):
    N = 1 << log_n
    moduli = pkg.primes_create(N, [bits] * L)
    ctx = pkg.Context(log_n, moduli)
    tabs = [O.Tables(log_n, q) for q in moduli]
    rng = np.random.default_rng(log_n * 7 + L)
    a_hat = np.stack([np.stack([rng.integers(0, q, N, dtype=np.uint64) for q in moduli]) for _ in range(n_poly)])
    a_hat[0, 0, 5] = np.uint64(2 ** 64 - 3)                     # words >= q are taken mod q
    a_hat[-1, -1, N - 1] += np.uint64(moduli[-1])
    red = np.stack([np.stack([a_hat[p, l] % np.uint64(moduli[l]) for l in range(L)]) for p in range(n_poly)])
    want = np.stack([O.ntt_inverse_batch(red[p].copy(), tabs) for p in range(n_poly)])
    out, verdict, sums = ctx.nwt_backward_protected_host(a_hat.copy())
    assert (out == want).all()
    assert not verdict.any() and (sums[:, 0] == sums[:, 1]).all()
    for l in (0, L - 1):
        w, w_hat = ctx.abft_weights(l)
        for p in (0, n_poly - 1):
            assert int(sums[p * L + l, 0]) == _dot_mod(w_hat, red[p, l], moduli[l])
            assert int(sums[p * L + l, 1]) == _dot_mod(w, want[p, l], moduli[l])
    # the unprotected entry gives the same bits, and forward-protected of the result closes the loop with the same checksum
    assert (ctx.nwt_backward_host(a_hat.copy()) == want).all()
    back, v2, s2 = ctx.nwt_forward_protected_host(want.copy())
    assert (back == red).all() and not v2.any() and (s2[:, 0] == sums[:, 1]).all()


def test_protected_inverse_rejects_limbs_without_inverse_tables(pkg):
    import ctypes as C
    B = pkg.binding
    N = 4096
    q = pkg.primes_create(N, [50])[0]
    rp, rps, irp, irps, ni, nis = pkg.host_tables(12, q)
    h = B.vp()
    B.check(B.lib().b200fhe_ctx_create_empty(12, 1, 0, C.byref(h)))
    B.check(B.lib().b200fhe_ctx_set_tables(h, 0, q, B._p(rp), B._p(rps), None, None, 0, 0, None))   # ntt_test.cu:65-66: nullptr inverse tables
    data = np.zeros(N, dtype=np.uint64)
    verdict = np.zeros(1, dtype=np.uint32)
    rc = B.lib().b200fhe_nwt_backward_protected_host(h, B._p(data), 1, 1, 0, verdict.ctypes.data_as(C.POINTER(C.c_uint32)), None)
    assert rc == -1 and b"inverse tables" in B.lib().b200fhe_last_error()
    B.lib().b200fhe_ctx_destroy(h)
