"""The encrypted dot product of reliability_test/dotprod_test.cu:95-185 end to end, with every evaluator step on the GPU
(multiply -> relinearize -> mod_switch_to_next -> elementwise check -> rotate-and-add over a row -> result[0] +
result[row_size]) and the roles the harness takes from Phantom's host classes (batch encoder, key generation,
encryption, decryption) played by test-side oracle arithmetic.  Fault-free the decrypted dot product is EXACT; with one
bit flipped in x's ciphertext (inject_bitflip_ciphertext, dotprod_test.cu:38-61) essentially every symbol is wrong,
the propagation law the reference's data shows (reliability_test/data)."""
import numpy as np
import pytest

from ks_common import _neg, _ntt_small, make_switch_key

pytestmark = pytest.mark.gpu

T = 12289                      # plain modulus: prime, = 1 mod 2N for N <= 2^11 (PlainModulus::Batching role)


def _dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).cuda()


def _host(t):
    return t.cpu().numpy().view(np.uint64)


def _signed_lift(v, big=1 << 62):
    return np.array([int(x) - big if int(x) > big // 2 else int(x) for x in v], dtype=object)


class Party:
    """secret key holder: encode / encrypt / decrypt / key generation on the host (test infrastructure)"""

    def __init__(self, O, rng, log_n, moduli, n_q, n_p, dnum):
        self.O, self.rng, self.log_n, self.N = O, rng, log_n, 1 << log_n
        self.moduli, self.n_q, self.n_p, self.dnum = moduli, n_q, n_p, dnum
        self.tabs = [O.Tables(log_n, q) for q in moduli]
        self.t_tab = O.Tables(log_n, T)
        self.s = rng.integers(-1, 2, self.N)
        self.s_h = _ntt_small(O, self.s, moduli, self.tabs)

    def encode(self, slots):                    # slot k = evaluation at psi_t^(2k+1), bit-reversed order as the NTT gives it
        return self.O.ntt_inverse(np.array(slots, dtype=np.uint64), self.t_tab)

    def decode(self, m):
        return self.O.ntt_forward(np.array(m, dtype=np.uint64), self.t_tab)

    def encrypt(self, m, seed):                 # (-a s + t e + m, a) over the n_q ciphertext moduli
        qs, tq = self.moduli[:self.n_q], self.tabs[:self.n_q]
        a = self.O.uniform_limbs(qs, self.N, seed=seed)
        me = _ntt_small(self.O, m.astype(np.int64) + T * np.rint(self.rng.normal(0, 3.2, self.N)).astype(np.int64), qs, tq)
        sh = self.s_h[:self.n_q]
        c0 = self.O.pmult_accum(np.stack([_neg(a, qs), me])[:, None], np.stack([sh, np.ones_like(sh)]), qs)[0]
        return np.stack([c0, a])

    def decrypt(self, ct, n_limbs):             # exact CRT lift of c0 + c1 s, centred, mod t
        qs, tq = self.moduli[:n_limbs], self.tabs[:n_limbs]
        sh = self.s_h[:n_limbs]
        d = self.O.ntt_inverse_batch(self.O.pmult_accum(ct[:, None], np.stack([np.ones_like(sh), sh]), qs)[0], tq)
        Q = 1
        for q in qs:
            Q *= int(q)
        x = np.zeros(self.N, dtype=object)
        for l, q in enumerate(qs):
            Ql = Q // int(q)
            x = (x + d[l].astype(object) * (pow(Ql, -1, int(q)) * Ql)) % Q
        x = np.where(x > Q // 2, x - Q, x)
        return np.array([int(v) % T for v in x], dtype=np.uint64)

    def switch_key_from(self, s_from, n_q):
        """key switching s_from -> s over the first n_q ciphertext moduli + the special primes"""
        idx = list(range(n_q)) + list(range(self.n_q, self.n_q + self.n_p))
        mods = [self.moduli[i] for i in idx]
        tabs = [self.tabs[i] for i in idx]
        return make_switch_key(self.O, self.rng, self.s, s_from, mods, n_q, self.n_p, self.dnum_at(n_q), tabs, t=T)

    def dnum_at(self, n_q):
        return self.dnum if n_q % self.dnum == 0 else 1


@pytest.mark.parametrize("flip", [None, (0, 1, 777, 13), (1, 0, 5, 59)])
def test_encrypted_dot_product_pipeline(pkg, O, flip):
    import torch
    from fhe_reliability_gpu_b200.evaluator import Evaluator
    log_n, n_q, n_p, dnum = 10, 4, 2, 2
    N = 1 << log_n
    moduli = pkg.primes_create(N, [60] * (n_q + n_p))
    rng = np.random.default_rng(2025)
    party = Party(O, rng, log_n, moduli, n_q, n_p, dnum)
    ctx = pkg.Context(log_n, moduli)                                          # level 0: q0..q3 + P
    q_ctx = pkg.Context(log_n, moduli[:n_q])                                  # for the mod-switch
    lvl1 = moduli[:n_q - 1] + moduli[n_q:]
    ctx1 = pkg.Context(log_n, lvl1)                                           # level 1: q0..q2 + P
    ev0 = Evaluator(ctx, n_q, n_p, dnum, plain_modulus=T)
    ev1 = Evaluator(ctx1, n_q - 1, n_p, party.dnum_at(n_q - 1), plain_modulus=T)

    # 1) inputs and CPU baseline (dotprod_test.cu:86-97)
    bits = 8
    x, y = rng.integers(0, 1 << bits, N), rng.integers(0, 1 << bits, N)
    baseline = (x * y) % T
    # 2-3) encode + encrypt
    cx, cy = party.encrypt(party.encode(x), 11), party.encrypt(party.encode(y), 22)
    d_x, d_y = _dev(torch, cx), _dev(torch, cy)
    # 4) fault injection into x's ciphertext: (poly, limb, coefficient, bit)
    if flip:
        p, l, i, bit = flip
        pkg.flip_bit(d_x.data_ptr(), (p * n_q + l) * N + i, bit)
    # 5) multiply, relinearize, mod-switch
    s2 = _signed_lift(O.ntt_inverse(np.array((party.s_h[0].astype(object) ** 2) % int(moduli[0]), dtype=np.uint64), party.tabs[0]),
                      int(moduli[0]))
    rk = party.switch_key_from(s2, n_q)
    relin = (_dev(torch, rk[:, :, :n_q]), _dev(torch, rk[:, :, n_q:]))
    xy = ev0.rescale_to_next(ev0.relinearize(ev0.multiply(d_x, d_y), relin), q_ctx)           # [2][n_q-1][N]
    torch.cuda.synchronize()
    # 6) decrypt + decode, compare elementwise (the mod-switch left the factor q_last^-1 on the plaintext)
    fix = int(moduli[n_q - 1]) % T
    prod = (party.decode(party.decrypt(_host(xy), n_q - 1)).astype(object) * fix) % T
    symbol_errors = int((prod != baseline.astype(object)).sum())
    if flip is None:
        assert symbol_errors == 0
    else:
        assert symbol_errors > 0.99 * N                       # symbol error rate ~1.0 (data/flipimpact_*.csv)
    # 7) rotate-and-add over a row: Galois elements 5^(2^i) (rotate_inplace by step = 2^i)
    step = 1
    while step < N // 2:
        g = pow(5, step, 2 * N)
        sg = _signed_lift(O.galois_coeff([int(v) % (1 << 62) for v in party.s], g, 1 << 62))
        gk = party.switch_key_from(sg, n_q - 1)
        key = (_dev(torch, gk[:, :, :n_q - 1]), _dev(torch, gk[:, :, n_q - 1:]))
        xy = ev1.add_inplace(xy, ev1.rotate_galois(xy, g, key))
        step <<= 1
    torch.cuda.synchronize()
    # 8) decrypt + decode: every slot of a row now holds the row's sum; result[0] + result[row_size]
    res = (party.decode(party.decrypt(_host(xy), n_q - 1)).astype(object) * fix) % T
    orbit, e = set(), 1
    for _ in range(N // 2):
        orbit.add(e)
        e = e * 5 % (2 * N)
    br = lambda v: int(format(v, "0%db" % log_n)[::-1], 2)
    row0 = [i for i in range(N) if (2 * br(i) + 1) in orbit]               # slots whose evaluation point is psi^(5^j)
    row1 = [i for i in range(N) if (2 * br(i) + 1) not in orbit]
    expected = int(baseline.sum()) % T
    if flip is None:
        assert len({int(res[i]) for i in row0}) == 1 and len({int(res[i]) for i in row1}) == 1
        assert int(res[row0[0]]) == int(sum(int(baseline[i]) for i in row0)) % T
        assert (int(res[row0[0]]) + int(res[row1[0]])) % T == expected      # "Dot product matches CPU result."
    else:
        assert (int(res[row0[0]]) + int(res[row1[0]])) % T != expected      # "MISMATCH detected!"
