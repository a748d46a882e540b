"""dotprod_test <bits_per_symbol> <num_symbols> -- the encrypted dot-product experiment of
reliability_test/dotprod_test.cu:67-228 on this library:

    python -m fhe_reliability_gpu_b200.dotprod <bits_per_symbol> <num_symbols> [--seed S] [--log-n 14]

Same parameters (BGV, N = 16384, six 50-bit primes of which two are special, 20-bit batching plain modulus,
dotprod_test.cu:201-206), same flow (encode, encrypt, inject bit flips into x's ciphertext on the device, multiply,
relinearize, mod_switch_to_next, elementwise comparison, rotate-and-add over a row, result[0] + result[row_size]) and
the same report lines.  Every homomorphic step runs on the GPU (evaluator.Evaluator); the roles of Phantom's host
classes (batch encoder, key generator, encryptor, decryptor) are played here with numpy for sampling / CRT and the
library's own NTT and multiply kernels for the polynomial arithmetic.  No CPU fallback: without the CUDA library or a
device this fails."""
import argparse
import sys

import numpy as np

from . import binding as B
from .evaluator import Evaluator

GEN = 3                                   # SEAL / Phantom Galois generator (GaloisTool::generator_)


def _bitrev(v, bits):
    return int(format(v, "0%db" % bits)[::-1], 2)


class Host:
    """Secret-key holder: the host classes dotprod_test.cu uses (PhantomSecretKey, PublicKey, BatchEncoder)."""

    def __init__(self, rng, log_n, moduli, n_q, n_p, dnum, t):
        self.rng, self.log_n, self.N = rng, log_n, 1 << log_n
        self.moduli, self.n_q, self.n_p, self.dnum, self.t = [int(q) for q in moduli], n_q, n_p, dnum, int(t)
        self.ctx = B.Context(log_n, self.moduli)
        self.ctx_t = B.Context(log_n, [self.t])
        self.s = rng.integers(-1, 2, self.N)
        self.s_h = self.ntt(self.s)
        N, two_n = self.N, 2 * self.N
        pos, self.index_map = 1, np.zeros(N, dtype=np.int64)          # BatchEncoder matrix_reps_index_map
        for i in range(N // 2):
            self.index_map[i] = _bitrev((pos - 1) >> 1, log_n)
            self.index_map[N // 2 + i] = _bitrev((two_n - pos - 1) >> 1, log_n)
            pos = pos * GEN % two_n

    # ---- polynomial helpers: residues [L][N] uint64, NTT form unless said otherwise; all products on the GPU
    def ntt(self, signed, limbs=None):
        limbs = list(range(len(self.moduli))) if limbs is None else limbs
        a = np.stack([np.array([int(v) % self.moduli[l] for v in signed], dtype=np.uint64) for l in limbs])
        out = np.empty_like(a)
        for i, l in enumerate(limbs):                               # one limb at a time: arbitrary limb subsets
            out[i] = self.ctx.nwt_forward_host(a[i].reshape(1, 1, self.N), L=1, start_mod=l).reshape(self.N)
        return out

    def mac(self, pairs, limbs):
        """sum of products of NTT-form polynomials given on `limbs`: pairs = [(x [len(limbs)][N], y ...), ...]"""
        out = np.empty((len(limbs), self.N), dtype=np.uint64)
        for i, l in enumerate(limbs):
            ct = np.stack([x[i] for x, _ in pairs]).reshape(len(pairs), 1, 1, self.N)
            pt = np.stack([y[i] for _, y in pairs]).reshape(len(pairs), 1, self.N)
            out[i] = self.ctx.pmult_accum_host(ct, pt, L=1, start_mod=l).reshape(self.N)
        return out

    def uniform(self, limbs):
        return np.stack([self.rng.integers(0, self.moduli[l], self.N, dtype=np.uint64) for l in limbs])

    def neg(self, a, limbs):
        q = np.array([self.moduli[l] for l in limbs], dtype=np.uint64)[:, None]
        return (q - a) % q

    def noise(self, limbs):
        return self.ntt(np.rint(self.rng.normal(0, 3.2, self.N)).astype(np.int64) * self.t, limbs)

    # ---- the Phantom host roles
    def encode(self, values):
        slots = np.zeros(self.N, dtype=np.uint64)
        slots[self.index_map] = np.array(values, dtype=np.uint64) % np.uint64(self.t)
        return self.ctx_t.nwt_backward_host(slots.reshape(1, 1, self.N)).reshape(self.N)

    def decode(self, m):
        slots = self.ctx_t.nwt_forward_host(np.array(m, dtype=np.uint64).reshape(1, 1, self.N)).reshape(self.N)
        return slots[self.index_map]

    def gen_publickey(self):
        limbs = list(range(self.n_q))
        a = self.uniform(limbs)
        ones = np.ones_like(a)
        self.pk = np.stack([self.mac([(self.neg(a, limbs), self.s_h[:self.n_q]), (self.noise(limbs), ones)], limbs), a])

    def encrypt_asymmetric(self, m):
        limbs = list(range(self.n_q))
        u = self.ntt(self.rng.integers(-1, 2, self.N), limbs)
        ones = np.ones_like(u)
        mh = self.ntt(m.astype(np.int64), limbs)
        c0 = self.mac([(self.pk[0], u), (self.noise(limbs), ones), (mh, ones)], limbs)
        c1 = self.mac([(self.pk[1], u), (self.noise(limbs), ones)], limbs)
        return np.stack([c0, c1])

    def decrypt(self, ct, n_limbs):
        limbs = list(range(n_limbs))
        d = self.mac([(ct[0], np.ones_like(ct[0])), (ct[1], self.s_h[:n_limbs])], limbs)
        coeff = np.stack([self.ctx.nwt_backward_host(d[i].reshape(1, 1, self.N), L=1, start_mod=l).reshape(self.N) for i, l in enumerate(limbs)])
        Q = 1
        for l in limbs:
            Q *= self.moduli[l]
        x = np.zeros(self.N, dtype=object)
        for i, l in enumerate(limbs):
            Ql = Q // self.moduli[l]
            x = (x + coeff[i].astype(object) * (pow(Ql, -1, self.moduli[l]) * Ql)) % Q
        x = np.where(x > Q // 2, x - Q, x)
        return np.array([int(v) % self.t for v in x], dtype=np.uint64)

    def switch_key(self, s_from_signed, n_q, dnum):
        """(evk_q [dnum][2][n_q][N], evk_p [dnum][2][n_p][N]): b_d = -a_d s + t e_d + P B_d s_from, a_d."""
        limbs = list(range(n_q)) + list(range(self.n_q, self.n_q + self.n_p))
        alpha = n_q // dnum
        P = 1
        for l in limbs[n_q:]:
            P *= self.moduli[l]
        sf, sh = self.ntt(s_from_signed, limbs), self.s_h[limbs]
        ones = np.ones_like(sh)
        evk = np.zeros((dnum, 2, len(limbs), self.N), dtype=np.uint64)
        for d in range(dnum):
            a = self.uniform(limbs)
            factor = np.zeros_like(a)
            for i in range(d * alpha, (d + 1) * alpha):
                factor[i, :] = P % self.moduli[limbs[i]]
            evk[d, 0] = self.mac([(self.neg(a, limbs), sh), (self.noise(limbs), ones), (factor, sf)], limbs)
            evk[d, 1] = a
        return np.ascontiguousarray(evk[:, :, :n_q]), np.ascontiguousarray(evk[:, :, n_q:])

    def galois_signed(self, g):
        out = np.zeros(self.N, dtype=np.int64)
        for i, v in enumerate(self.s):
            idx = (i * g) % (2 * self.N)
            if idx >= self.N:
                out[idx - self.N] = -v
            else:
                out[idx] = v
        return out

    def square_signed(self):
        q0 = self.moduli[0]
        sq = self.mac([(self.s_h[:1], self.s_h[:1])], [0])
        c = self.ctx.nwt_backward_host(sq.reshape(1, 1, self.N), L=1, start_mod=0).reshape(self.N)
        return np.array([int(v) - q0 if int(v) > q0 // 2 else int(v) for v in c], dtype=object)


def _fmt(v):
    v = [int(x) for x in v]
    return "[ " + ", ".join(str(x) for x in v[:3]) + ", ..., " + ", ".join(str(x) for x in v[-3:]) + " ]"


def run(bits_per_symbol, num_symbols, seed=None, log_n=14, prime_bits=50, out=sys.stdout, err=sys.stderr):
    """Returns a dict with the numbers of the report (symbol_errors, bit_errors, result, expected, match)."""
    import torch
    if not torch.cuda.is_available():
        raise B.B200FHEError("dotprod needs a CUDA device (no CPU fallback)")
    rng = np.random.default_rng(seed)
    N = 1 << log_n
    n_q, n_p = 4, 2                                            # {50 x 6}, special_modulus_size 2 (dotprod_test.cu:204-205)
    dnum = 2
    moduli = B.primes_create(N, [prime_bits] * (n_q + n_p))
    t = B.primes_create(N, [20])[0]                            # PlainModulus::Batching(N, 20)
    host = Host(rng, log_n, moduli, n_q, n_p, dnum, t)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).cuda()
    back = lambda x: x.cpu().numpy().view(np.uint64)
    ev0 = Evaluator(host.ctx, n_q, n_p, dnum, plain_modulus=t)
    q_ctx = B.Context(log_n, moduli[:n_q])
    ctx1 = B.Context(log_n, moduli[:n_q - 1] + moduli[n_q:])
    ev1 = Evaluator(ctx1, n_q - 1, n_p, n_q - 1, plain_modulus=t)    # level 1: three primes, one prime per digit

    print("Example: BGV HomMul test", file=out)
    host.gen_publickey()
    rk = host.switch_key(host.square_signed(), n_q, dnum)
    relin = (dev(rk[0]), dev(rk[1]))
    slot_count, row_size = N, N // 2
    print("Plaintext matrix row size: %d" % row_size, file=out)
    in1, in2 = rng.integers(0, 100, slot_count), rng.integers(0, 100, slot_count)
    print("Input vector 1: " + _fmt(in1), file=out)
    print("Input vector 2: " + _fmt(in2), file=out)
    baseline = (in1 * in2) % t
    d_x, d_y = dev(host.encrypt_asymmetric(host.encode(in1))), dev(host.encrypt_asymmetric(host.encode(in2)))
    total = 2 * n_q * N                                        # inject_bitflip_ciphertext, dotprod_test.cu:38-61
    for _ in range(num_symbols):
        idx = int(rng.integers(0, total))
        for _ in range(bits_per_symbol):
            bit = int(rng.integers(0, 64))
            B.flip_bit(d_x.data_ptr(), idx, bit)
            torch.cuda.synchronize()
            print("Injected bitflip @ idx=%d, bit=%d" % (idx, bit), file=err)
    print("Compute x * y homomorphically...", file=out)
    xy = ev0.rescale_to_next(ev0.relinearize(ev0.multiply(d_x, d_y), relin), q_ctx)
    torch.cuda.synchronize()
    fix = moduli[n_q - 1] % t                                  # mod_switch_to_next leaves m q_last^-1: BGV's correction factor
    prod = (host.decode(host.decrypt(back(xy), n_q - 1)).astype(object) * fix) % t
    wrong = prod != baseline.astype(object)
    symbol_errors = int(wrong.sum())
    bit_errors = sum(bin(int(a) ^ int(b)).count("1") for a, b in zip(baseline[wrong], prod[wrong]))
    print("Raw product vector: " + _fmt(prod), file=out)
    print("CPU baseline      : " + _fmt(baseline), file=out)
    print("Elementwise symbol errors: %d / %d" % (symbol_errors, slot_count), file=out)
    print("Elementwise Hamming distance (bit errors): %d" % bit_errors, file=out)
    step = 1
    while step < row_size:                                     # rotate-and-add, dotprod_test.cu:143-149
        g = pow(GEN, step, 2 * N)
        gk = host.switch_key(host.galois_signed(g), n_q - 1, n_q - 1)
        xy = ev1.add_inplace(xy, ev1.rotate_galois(xy, g, (dev(gk[0]), dev(gk[1]))))
        step <<= 1
    torch.cuda.synchronize()
    res = (host.decode(host.decrypt(back(xy), n_q - 1)).astype(object) * fix) % t
    result_full = (int(res[0]) + int(res[row_size])) % t
    expected = int(sum(int(v) for v in baseline)) % t
    dp_bits = bin(result_full ^ expected).count("1")
    print("Decrypted dot product = %d" % result_full, file=out)
    print("Expected (CPU)         = %d" % expected, file=out)
    print("Dot product bit errors (Hamming distance): %d" % dp_bits, file=out)
    abs_diff = abs(result_full - expected)
    print("Absolute difference   = %d" % abs_diff, file=out)
    if expected:
        print("Percentage error      = %g%%" % (100.0 * abs_diff / expected), file=out)
    else:
        print("Percentage error      = undefined (expected is zero)", file=out)
    print("✔ Dot product matches CPU result." if result_full == expected else "✖ MISMATCH detected!", file=out)
    return {"symbol_errors": symbol_errors, "bit_errors": bit_errors, "slot_count": slot_count, "result": result_full,
            "expected": expected, "match": result_full == expected}


def main(argv=None):
    argv = sys.argv[1:] if argv is None else argv
    ap = argparse.ArgumentParser(prog="dotprod_test", usage="%(prog)s <bits_per_symbol> <num_symbols>")
    ap.add_argument("bits_per_symbol", type=int)
    ap.add_argument("num_symbols", type=int)
    ap.add_argument("--seed", type=int, default=None)
    ap.add_argument("--log-n", type=int, default=14)
    try:
        a = ap.parse_args(argv)
    except SystemExit:
        return 1                                               # dotprod_test.cu:190-194: usage on stderr, exit 1
    run(a.bits_per_symbol, a.num_symbols, seed=a.seed, log_n=a.log_n)
    return 0


if __name__ == "__main__":
    sys.exit(main())
