"""CPU: the oracle's key-switch composition (oracle.keyswitch_*) really is a key-switch -- with an RLWE switching
key, out0 + out1 s - c s' is one small integer polynomial -- so the GPU parity tests compare against the right thing."""
import numpy as np
import pytest

from ks_common import check_switch_noise, make_switch_key


@pytest.mark.parametrize("log_n,n_q,n_p,dnum", [(8, 6, 2, 3), (9, 4, 4, 1), (8, 6, 1, 6)])
def test_oracle_keyswitch_is_a_keyswitch(O, log_n, n_q, n_p, dnum):
    N = 1 << log_n
    moduli = O.primes_create(N, [50] * (n_q + n_p))
    tabs = [O.Tables(log_n, q) for q in moduli]
    rng = np.random.default_rng(log_n + dnum)
    s, s_new = rng.integers(-1, 2, N), rng.integers(-1, 2, N)
    evk = make_switch_key(O, rng, s, s_new, moduli, n_q, n_p, dnum, tabs)
    c = O.uniform_limbs(moduli[:n_q], N, seed=3)
    out = O.keyswitch_hybrid(c, evk, moduli, n_q, n_p, dnum, tabs)
    noise = check_switch_noise(O, out, c, s, s_new, moduli[:n_q], tabs[:n_q], bound_bits=30)
    assert any(int(v) != 0 for v in noise)            # a real key adds real noise


def test_oracle_modup_raises_digit_exactly_up_to_small_multiple(O):
    """up_d on a target limb = (x_d + e D_d) mod q with 0 <= e < alpha, where x_d = CRT of the digit's residues."""
    log_n, n_q, n_p, dnum = 6, 4, 2, 2
    N = 1 << log_n
    moduli = O.primes_create(N, [40] * (n_q + n_p))
    tabs = [O.Tables(log_n, q) for q in moduli]
    c = O.uniform_limbs(moduli[:n_q], N, seed=11)
    up = O.keyswitch_modup(c, moduli, n_q, n_p, dnum, tabs)
    x = O.ntt_inverse_batch(c, tabs[:n_q])
    alpha = n_q // dnum
    for d in range(dnum):
        dig = list(range(d * alpha, (d + 1) * alpha))
        D = 1
        for g in dig:
            D *= int(moduli[g])
        coeff = np.stack([O.ntt_inverse(up[d, g], tabs[g]) for g in range(n_q + n_p)])
        for i in range(0, N, 7):
            xi = 0
            for g in dig:
                Dg = D // int(moduli[g])
                xi = (xi + int(x[g, i]) * pow(Dg, -1, int(moduli[g])) % int(moduli[g]) * Dg) % D
            es = {e for e in range(alpha) if all((xi + e * D) % int(moduli[g]) == int(coeff[g, i]) for g in range(n_q + n_p))}
            assert len(es) == 1


def test_oracle_bgv_keyswitch_keeps_the_plaintext_modulus(O):
    """BGV flavour (t-scaled ModDown, keys with noise t e_d): out0 + out1 s - c s' is small AND a multiple of t."""
    log_n, n_q, n_p, dnum, t = 8, 6, 2, 3, 65537
    N = 1 << log_n
    moduli = O.primes_create(N, [50] * (n_q + n_p))
    tabs = [O.Tables(log_n, q) for q in moduli]
    rng = np.random.default_rng(12)
    s, s_new = rng.integers(-1, 2, N), rng.integers(-1, 2, N)
    evk = make_switch_key(O, rng, s, s_new, moduli, n_q, n_p, dnum, tabs, t=t)
    c = O.uniform_limbs(moduli[:n_q], N, seed=3)
    noise = check_switch_noise(O, O.keyswitch_hybrid(c, evk, moduli, n_q, n_p, dnum, tabs, t=t), c, s, s_new, moduli[:n_q], tabs[:n_q], 40)
    assert not (noise % t).any() and np.abs(noise).max() > 0
    plain = check_switch_noise(O, O.keyswitch_hybrid(c, evk, moduli, n_q, n_p, dnum, tabs), c, s, s_new, moduli[:n_q], tabs[:n_q], 40)
    assert (plain % t).any()                      # without the correction the residual is not a multiple of t
