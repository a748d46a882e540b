"""GPU parity of the evaluator steps around the key-switch (arithmetic only; call sites
reliability_test/dotprod_test.cu:113-152): tensor product, add/sub, Galois permutation and rescale bit-exact against
the oracle, and the two compositions the dot-product pipeline is made of -- multiply + relinearize, rotate -- checked
by decrypting with the secret key (the residual must be one small integer polynomial)."""
import numpy as np
import pytest

from ks_common import _neg, _ntt_small, make_switch_key

pytestmark = pytest.mark.gpu


def _dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).cuda()


def _host(t):
    return t.cpu().numpy().view(np.uint64)


def _setup(pkg, O, log_n, n_q, n_p, dnum):
    from fhe_reliability_gpu_b200.evaluator import Evaluator
    N = 1 << log_n
    moduli = pkg.primes_create(N, [60] * (n_q + n_p))
    ctx = pkg.Context(log_n, moduli)
    tabs = [O.Tables(log_n, q) for q in moduli]
    return N, moduli, ctx, tabs, Evaluator(ctx, n_q, n_p, dnum)


def _encrypt(O, rng, m, s_h, moduli, tabs, seed):
    """(c0, c1) = (-a s + m + e, a), NTT form over `moduli`"""
    N = len(m)
    a = O.uniform_limbs(moduli, N, seed=seed)
    me = _ntt_small(O, np.asarray(m) + np.rint(rng.normal(0, 3.2, N)).astype(np.int64), moduli, tabs)
    c0 = O.pmult_accum(np.stack([_neg(a, moduli), me])[:, None], np.stack([s_h, np.ones_like(s_h)]), moduli)[0]
    return np.stack([c0, a])


def _residual(O, ct_terms, pt_terms, qs, tq, bound_bits):
    r = O.pmult_accum(np.stack(ct_terms)[:, None], np.stack(pt_terms), qs)[0]
    r = O.ntt_inverse_batch(r, tq).astype(np.int64)
    q = np.array([int(v) for v in qs], dtype=np.int64)[:, None]
    r = np.where(r > q // 2, r - q, r)
    assert int(np.abs(r).max()) < 1 << bound_bits, "residual is not small: 2^%.1f" % np.log2(float(np.abs(r).max()))
    assert (r == r[0]).all(), "residual differs between limbs"
    return r[0]


@pytest.mark.parametrize("log_n,L", [(4, 2), (12, 5), (14, 3)])
def test_tensor_add_sub_vs_oracle(pkg, O, log_n, L):
    import torch
    N, moduli, ctx, tabs, ev = _setup(pkg, O, log_n, L, 1, 1)
    a = np.stack([O.uniform_limbs(moduli[:L], N, seed=1), O.uniform_limbs(moduli[:L], N, seed=2)])
    b = np.stack([O.uniform_limbs(moduli[:L], N, seed=3), O.uniform_limbs(moduli[:L], N, seed=4)])
    a[0, 0, :3] = np.uint64(2 ** 64 - 1)                        # words >= q are taken mod q
    got = _host(ev.multiply(_dev(torch, a), _dev(torch, b)))
    assert np.array_equal(got, O.ct_tensor(a, b, moduli[:L]))
    a[0, 0, :3] = 0
    q = np.array(moduli[:L], dtype=object)[None, :, None]
    s = _host(ev.add_inplace(_dev(torch, a), _dev(torch, b)))
    assert np.array_equal(s, np.array((a.astype(object) + b.astype(object)) % q, dtype=np.uint64))
    d = _host(ev.sub_inplace(_dev(torch, a), _dev(torch, b)))
    assert np.array_equal(d, np.array((a.astype(object) - b.astype(object)) % q, dtype=np.uint64))


@pytest.mark.parametrize("log_n", [3, 10, 13])
def test_galois_ntt_vs_coefficient_domain_automorphism(pkg, O, log_n):
    import torch
    N, moduli, ctx, tabs, ev = _setup(pkg, O, log_n, 2, 1, 1)
    coeff = O.uniform_limbs(moduli[:2], N, seed=log_n)
    ntt = np.stack([O.ntt_forward(coeff[l], tabs[l]) for l in range(2)])
    ct = _dev(torch, np.stack([ntt, ntt[::-1] * 0 + ntt]))    # two polynomials
    for g in sorted({3, 5, 25 % (2 * N), 2 * N - 1, pow(5, N // 4 + 1, 2 * N)}):
        got = _host(ev.apply_galois(ct, g))
        want = np.stack([O.ntt_forward(O.galois_coeff(coeff[l], g, moduli[l]), tabs[l]) for l in range(2)])
        assert np.array_equal(got[0], want) and np.array_equal(got[1], want), g
    with pytest.raises(pkg.B200FHEError):
        ev.apply_galois(ct, 4)                                  # even elements are not automorphisms


def test_rescale_is_exact_floor_division(pkg, O):
    """rescale_to_next == floor(X / q_last) limb by limb, for an integer polynomial X given by its residues."""
    import torch
    log_n, L = 6, 4
    N, moduli, ctx, tabs, ev = _setup(pkg, O, log_n, L, 1, 1)
    q_ctx = pkg.Context(log_n, moduli[:L])
    Q = 1
    for q in moduli[:L]:
        Q *= int(q)
    rng = np.random.default_rng(5)
    X = [[int.from_bytes(rng.bytes(30), "little") % Q for _ in range(N)] for _ in range(2)]
    ct = np.stack([np.stack([O.ntt_forward(np.array([x % int(q) for x in Xp], dtype=np.uint64), t) for q, t in zip(moduli[:L], tabs)])
                   for Xp in X])
    got = _host(ev.rescale_to_next(_dev(torch, ct), q_ctx))
    for p in range(2):
        for l in range(L - 1):
            coeff = O.ntt_inverse(got[p, l], tabs[l])
            assert [int(v) for v in coeff] == [(x // int(moduli[L - 1])) % int(moduli[l]) for x in X[p]]
        assert np.array_equal(got[p], O.keyswitch_moddown(ct[p], moduli[:L], L - 1, 1, tabs[:L]))


@pytest.mark.parametrize("log_n,n_q,n_p,dnum", [(10, 4, 2, 2), (12, 6, 2, 3)])
def test_multiply_relinearize_decrypts_to_the_product(pkg, O, log_n, n_q, n_p, dnum):
    import torch
    N, moduli, ctx, tabs, ev = _setup(pkg, O, log_n, n_q, n_p, dnum)
    qs, tq = moduli[:n_q], tabs[:n_q]
    rng = np.random.default_rng(log_n)
    s = rng.integers(-1, 2, N)
    s_h_all = _ntt_small(O, s, moduli, tabs)
    s_h = s_h_all[:n_q]
    # s^2 as a signed coefficient vector (|coeff| <= N): square in the NTT domain of one modulus, lift centred
    q0 = int(moduli[0])
    sq = O.ntt_inverse(np.array((s_h[0].astype(object) ** 2) % q0, dtype=np.uint64), tabs[0]).astype(object)
    s2 = np.array([int(v) - q0 if int(v) > q0 // 2 else int(v) for v in sq], dtype=object)
    relin = make_switch_key(O, rng, s, s2, moduli, n_q, n_p, dnum, tabs)
    key = (_dev(torch, relin[:, :, :n_q]), _dev(torch, relin[:, :, n_q:]))
    m1, m2 = rng.integers(-500, 500, N), rng.integers(-500, 500, N)
    ct1, ct2 = _encrypt(O, rng, m1, s_h, qs, tq, 101), _encrypt(O, rng, m2, s_h, qs, tq, 202)
    out = _host(ev.relinearize(ev.multiply(_dev(torch, ct1), _dev(torch, ct2)), key))
    ones = np.ones_like(s_h)
    dec1 = O.pmult_accum(ct1[:, None], np.stack([ones, s_h]), qs)[0]
    dec2 = O.pmult_accum(ct2[:, None], np.stack([ones, s_h]), qs)[0]
    noise = _residual(O, [out[0], out[1], _neg(dec1, qs)], [ones, s_h, dec2], qs, tq, 34)
    assert np.abs(noise).max() > 0
    # and the decryption really is m1 m2 + small: compare with the plaintext product up to the encryption noise
    m1h, m2h = _ntt_small(O, m1, qs, tq), _ntt_small(O, m2, qs, tq)
    _residual(O, [out[0], out[1], _neg(m1h, qs)], [ones, s_h, m2h], qs, tq, 34)


@pytest.mark.parametrize("log_n,n_q,n_p,dnum,g", [(10, 4, 2, 2, 3), (12, 6, 2, 3, 5), (11, 4, 4, 1, 2 * 2048 - 1)])
def test_rotate_decrypts_to_the_automorphed_plaintext(pkg, O, log_n, n_q, n_p, dnum, g):
    import torch
    N, moduli, ctx, tabs, ev = _setup(pkg, O, log_n, n_q, n_p, dnum)
    qs, tq = moduli[:n_q], tabs[:n_q]
    rng = np.random.default_rng(log_n + g)
    s = rng.integers(-1, 2, N)
    s_h = _ntt_small(O, s, qs, tq)
    big = 1 << 62
    sg = np.array([int(v) - big if int(v) > big // 2 else int(v) for v in O.galois_coeff([int(v) % big for v in s], g, big)], dtype=object)
    gkey = make_switch_key(O, rng, s, sg, moduli, n_q, n_p, dnum, tabs)          # from s(x^g) to s
    key = (_dev(torch, gkey[:, :, :n_q]), _dev(torch, gkey[:, :, n_q:]))
    m = rng.integers(-1000, 1000, N)
    ct = _encrypt(O, rng, m, s_h, qs, tq, 7)
    out = _host(ev.rotate_galois(_dev(torch, ct), g, key))
    ones = np.ones_like(s_h)
    dec = O.ntt_inverse_batch(O.pmult_accum(ct[:, None], np.stack([ones, s_h]), qs)[0], tq)        # m + e, coefficient domain
    dec_g = np.stack([O.ntt_forward(O.galois_coeff(dec[l], g, int(qs[l])), tq[l]) for l in range(n_q)])
    _residual(O, [out[0], out[1], _neg(dec_g, qs)], [ones, s_h, ones], qs, tq, 34)


def test_bgv_mod_switch_keeps_the_plaintext_up_to_the_correction_factor(pkg, O):
    """BGV: ct = (-a s + t e + m, a); after mod_switch_to_next (rescale with plain_modulus = t) the ciphertext decrypts
    EXACTLY to m q_last^-1 mod t -- the correction factor BGV libraries track (Phantom / SEAL mod_switch_to_next)."""
    import torch
    from fhe_reliability_gpu_b200.evaluator import Evaluator
    log_n, L, t = 10, 4, 65537
    N = 1 << log_n
    moduli = pkg.primes_create(N, [60] * (L + 1))
    ctx = pkg.Context(log_n, moduli)
    q_ctx = pkg.Context(log_n, moduli[:L])
    tabs = [O.Tables(log_n, q) for q in moduli]
    ev = Evaluator(ctx, L, 1, 1, plain_modulus=t)
    qs, tq = moduli[:L], tabs[:L]
    rng = np.random.default_rng(2)
    s = rng.integers(-1, 2, N)
    s_h = _ntt_small(O, s, qs, tq)
    m = rng.integers(0, t, N)
    a = O.uniform_limbs(qs, N, seed=5)
    me = _ntt_small(O, np.asarray(m) + t * np.rint(rng.normal(0, 3.2, N)).astype(np.int64), qs, tq)
    c0 = O.pmult_accum(np.stack([_neg(a, qs), me])[:, None], np.stack([s_h, np.ones_like(s_h)]), qs)[0]
    out = _host(ev.rescale_to_next(_dev(torch, np.stack([c0, a])), q_ctx))
    lo, tl = qs[:L - 1], tq[:L - 1]
    dec = O.ntt_inverse_batch(O.pmult_accum(out[:, None], np.stack([np.ones_like(s_h[:L - 1]), s_h[:L - 1]]), lo)[0], tl).astype(np.int64)
    q = np.array([int(v) for v in lo], dtype=np.int64)[:, None]
    dec = np.where(dec > q // 2, dec - q, dec)
    assert (dec == dec[0]).all() and np.abs(dec).max() < 1 << 40
    want = (m.astype(object) * pow(int(moduli[L - 1]), -1, t)) % t
    assert [int(v) % t for v in dec[0]] == [int(v) for v in want]
    # bit-exact against the oracle's t-scaled ModDown
    for p, poly in enumerate((c0, a)):
        assert np.array_equal(out[p], O.keyswitch_moddown(poly, qs, L - 1, 1, tq, t=t))
