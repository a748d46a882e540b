"""Shared helpers of the key-switch tests: RLWE switching-key generation and the noise check (test-side only)."""
import numpy as np


def _ntt_small(O, poly, moduli, tabs):
    """signed small-coefficient polynomial -> NTT domain over the given moduli"""
    return np.stack([O.ntt_forward(np.array([int(v) % q for v in poly], dtype=np.uint64), t) for q, t in zip(moduli, tabs)])


def _neg(a, moduli):
    q = np.array([int(v) for v in moduli], dtype=np.uint64)[:, None]
    return (q - a) % q


def make_switch_key(O, rng, s, s_new, moduli, n_q, n_p, dnum, tabs, sigma=3.2, t=1):
    """evk_d = (-a_d s + e_d + P B_d s_new, a_d) over Q u P, NTT domain; B_d = 1 mod the digit's moduli, 0 mod the rest of Q.
    Modular products through the oracle's pmult (C), so full-size keys take seconds.  t > 1: BGV keys (noise term t e_d)."""
    N = len(s)
    alpha = n_q // dnum
    P = 1
    for p in moduli[n_q:]:
        P *= int(p)
    s_h, sn_h = _ntt_small(O, s, moduli, tabs), _ntt_small(O, s_new, moduli, tabs)
    ones = np.ones_like(s_h)
    evk = np.zeros((dnum, 2, n_q + n_p, N), dtype=np.uint64)
    for d in range(dnum):
        a = O.uniform_limbs(moduli, N, seed=int(rng.integers(1 << 30)))
        e = _ntt_small(O, np.rint(rng.normal(0, sigma, N)).astype(np.int64) * int(t), moduli, tabs)
        factor = np.zeros_like(a)                                   # P B_d mod q_l (0 mod every p_j as well)
        for l in range(d * alpha, (d + 1) * alpha):
            factor[l, :] = P % int(moduli[l])
        ct = np.stack([_neg(a, moduli), e, factor])[:, None]        # [3 terms][1][L][N]
        evk[d, 0] = O.pmult_accum(ct, np.stack([s_h, ones, sn_h]), moduli)[0]
        evk[d, 1] = a
    return evk


def check_switch_noise(O, out, c, s, s_new, qs, tq, bound_bits=34):
    """out [2][n_q][N], c [n_q][N] (NTT domain): out0 + out1 s - c s_new must be ONE small integer polynomial."""
    s_h, sn_h = _ntt_small(O, s, qs, tq), _ntt_small(O, s_new, qs, tq)
    ct = np.stack([out[0], out[1], _neg(np.ascontiguousarray(c), qs)])[:, None]
    r = O.pmult_accum(ct, np.stack([np.ones_like(s_h), s_h, sn_h]), qs)[0]
    r = O.ntt_inverse_batch(r, tq).astype(np.int64)
    q = np.array([int(v) for v in qs], dtype=np.int64)[:, None]
    r = np.where(r > q // 2, r - q, r)
    assert int(np.abs(r).max()) < 1 << bound_bits, "key-switch noise is not small"
    assert (r == r[0]).all(), "noise differs between limbs (result is not one integer polynomial)"
    return r[0]
