"""The reference's Python prototype interfaces, served by the CUDA library.

Same names, argument meaning and result layout (Python lists of ints, natural order) as
  rfhe_framewk/src/negaclic_ntt.py:38  ntt(a, root, mod)          cyclic, natural order out
  rfhe_framewk/src/negaclic_ntt.py:77  intt(a, root, mod)
  rfhe_framewk/src/negaclic_ntt.py:86  negacyclic_ntt(a, psi, mod)
  rfhe_framewk/src/negaclic_ntt.py:102 negacyclic_intt(A, psi, mod)
  rfhe_framewk/src/negaclic_ntt.py:130 ecc_check(a, psi, mod) -> (checksum, checksum_hat, equal)
  reliability_test/four_step_ntt_prot.py:71 four_step_ntt(a, N, mod, g)
so a script swaps its import and keeps its call sites.

Every transform, twist multiply and checksum runs on the GPU through include/b200fhe.h.  The host
side only builds tables (root powers, the psi^-j twist vector: the same kind of precomputation
ntt_test.cu:59-69 does on the host) and reorders indices between the prototypes' natural order
and the library's SEAL order (out[bitrev(k)]).  Requirements inherited from the library: mod a
prime below 2^60, len(a) = 2^k with 2*len(a) dividing mod-1.  No CPU fallback: without the CUDA
library or a device every call raises B200FHEError.
"""
import numpy as np

from . import binding as B

MOD = 998244353      # four_step_ntt_prot.py:16-17
G = 3

_ctx_cache = {}


def _log2(n):
    k = n.bit_length() - 1
    if n < 2 or (1 << k) != n:
        raise ValueError("length must be a power of two >= 2")
    return k


def _bitrev_perm(k):
    idx = np.arange(1 << k, dtype=np.int64)
    r = np.zeros_like(idx)
    for b in range(k):
        r |= ((idx >> b) & 1) << (k - 1 - b)
    return r


def _ctx(n, psi, mod):
    key = (n, psi % mod, mod)
    c = _ctx_cache.get(key)
    if c is None:
        if len(_ctx_cache) > 16:
            _ctx_cache.clear()
        c = _ctx_cache[key] = B.Context.from_psi(_log2(n), [mod], [psi % mod])
    return c


def _arr(a, mod):
    return np.array([int(x) % mod for x in a], dtype=np.uint64)


def negacyclic_ntt(a, psi, mod):
    """out[k] = sum_j a_j psi^((2k+1) j) mod `mod`, natural order (negaclic_ntt.py:86-92)."""
    n = len(a)
    c = _ctx(n, psi, mod)
    out = c.nwt_forward_host(_arr(a, mod).reshape(1, 1, n)).reshape(n)
    return [int(x) for x in out[_bitrev_perm(_log2(n))]]


def negacyclic_intt(A, psi, mod):
    """Inverse of negacyclic_ntt (negaclic_ntt.py:102-109)."""
    n = len(A)
    c = _ctx(n, psi, mod)
    x = np.empty(n, dtype=np.uint64)
    x[_bitrev_perm(_log2(n))] = _arr(A, mod)          # natural order -> SEAL (bit-reversed) order
    return [int(v) for v in c.nwt_backward_host(x.reshape(1, 1, n)).reshape(n)]


def _sqrt_mod(a, p):
    """A square root of a modulo the prime p (Tonelli-Shanks); host table construction only."""
    a %= p
    if a == 0:
        return 0
    if pow(a, (p - 1) // 2, p) != 1:
        raise ValueError("root has no square root: 2*len(a) must divide mod-1")
    q, s = p - 1, 0
    while q % 2 == 0:
        q //= 2
        s += 1
    z = 2
    while pow(z, (p - 1) // 2, p) != p - 1:
        z += 1
    m, c, t, r = s, pow(z, q, p), pow(a, q, p), pow(a, (q + 1) // 2, p)
    while t != 1:
        i, t2 = 0, t
        while t2 != 1:
            t2 = t2 * t2 % p
            i += 1
        b = pow(c, 1 << (m - i - 1), p)
        m, c, t, r = i, b * b % p, t * b * b % p, r * b % p
    return r


def _twist(c, vals, base, mod):
    """vals[j] * base^j mod `mod`, multiplied on the GPU (b200fhe_pmult_accum, one term)."""
    n = len(vals)
    pw = np.empty(n, dtype=np.uint64)
    x = 1
    for j in range(n):
        pw[j] = x
        x = x * base % mod
    return c.pmult_accum_host(vals.reshape(1, 1, 1, n), pw.reshape(1, 1, n)).reshape(n)


def _psi_for(root, n, mod):
    psi = _sqrt_mod(root, mod)
    if pow(psi, n, mod) != mod - 1:          # the other square root is the primitive 2n-th root
        psi = mod - psi
    if pow(psi, n, mod) != mod - 1:
        raise ValueError("root is not a primitive len(a)-th root of unity")
    return psi


def ntt(a, root, mod):
    """Cyclic NTT, natural order out: X[k] = sum_j a_j root^(jk) (negaclic_ntt.py:38-57)."""
    n = len(a)
    psi = _psi_for(root % mod, n, mod)
    c = _ctx(n, psi, mod)
    b = _twist(c, _arr(a, mod), pow(psi, mod - 2, mod), mod)       # b_j = a_j psi^-j
    out = c.nwt_forward_host(b.reshape(1, 1, n)).reshape(n)
    return [int(x) for x in out[_bitrev_perm(_log2(n))]]


def intt(a, root, mod):
    """Inverse cyclic NTT scaled by n^-1 (negaclic_ntt.py:77-83)."""
    n = len(a)
    psi = _psi_for(root % mod, n, mod)
    c = _ctx(n, psi, mod)
    x = np.empty(n, dtype=np.uint64)
    x[_bitrev_perm(_log2(n))] = _arr(a, mod)
    b = c.nwt_backward_host(x.reshape(1, 1, n)).reshape(n)
    return [int(v) for v in _twist(c, b, psi, mod)]                # a_j = b_j psi^j


def four_step_ntt(a, N, mod=MOD, g=G):
    """reliability_test/four_step_ntt_prot.py:71-109: y_k = sum_t a_t w^(kt), w = g^((mod-1)/N)."""
    if len(a) != N:
        raise ValueError("len(a) != N")
    return ntt(a, pow(g, (mod - 1) // N, mod), mod)


def ntt_direct(a, N, mod=MOD, g=G):
    """four_step_ntt_prot.py:49-58 computes the same transform by the O(N^2) definition."""
    return four_step_ntt(a, N, mod, g)


def generate_weights(length):
    """negaclic_ntt.py:8-13 (the input-side ABFT weights the kernels regenerate on the fly)."""
    p = int(round(length ** 0.5))
    if p * p != length:
        raise AssertionError("Length must be a perfect square")
    return [(i % p) + (i // p) + 2 for i in range(length)]


def ecc_check(a, psi, mod, plan=()):
    """negaclic_ntt.py:130-149: (sum_i w_i a_i, sum_k w^_k a^_k, equal) with both sums accumulated
    inside the protected forward transform (b200fhe_nwt_forward_protected).  `plan` optionally
    injects faults (binding.make_plan entries) in place of the prototype's random input flips."""
    n = len(a)
    generate_weights(n)                                # same perfect-square requirement
    c = _ctx(n, psi, mod)
    _, verdict, sums = c.nwt_forward_protected_host(_arr(a, mod).reshape(1, 1, n), plan=plan)
    return int(sums[0, 0]), int(sums[0, 1]), not bool(verdict[0])


def bConv_ecc(residues, p_moduli, q_moduli, inject=()):
    """rfhe_framewk/src/baseConv.py:42-98: base conversion that carries a checksum and raises
    ValueError("ECC mismatch ...") when it does not survive.  residues[j][i] mod p_moduli[j] -> result[i][k] mod
    q_moduli[k] (the prototype's orientation).  Values are the HPS conversion's (SEAL / Phantom fast_convert_array:
    x + e P for a small e); the prototype's own bConv differs from it by multiples of P.
    Needs len(residues[0]) = 2^k and every modulus NTT-friendly for that length (the library's context rule)."""
    n = len(residues[0])
    mods = [int(v) for v in list(p_moduli) + list(q_moduli)]
    key = ("bconv", n, tuple(mods))
    c = _ctx_cache.get(key)
    if c is None:
        ctx = B.Context(_log2(n), mods)
        c = _ctx_cache[key] = (ctx, B.BaseConv(ctx, list(range(len(p_moduli))), list(range(len(p_moduli), len(mods)))))
    x = np.array([[int(v) % int(p) for v in row] for row, p in zip(residues, p_moduli)], dtype=np.uint64)
    out, verdict, _ = c[1].modup_protected_host(x, inject=inject)
    bad = np.nonzero(verdict)[0]
    if len(bad):
        i = int(bad[0])
        where = "source modulus index %d" % i if i < len(p_moduli) else "target modulus index %d" % (i - len(p_moduli))
        raise ValueError("ECC mismatch at %s" % where)
    return [[int(out[k][i]) for k in range(len(q_moduli))] for i in range(n)]


def four_step_with_protection_vector(a, N, mod=MOD, g=G, fault=None):
    """reliability_test/four_step_ntt_prot.py:196-252: returns (y, checks) with
    checks = {"stage1": {lhs, rhs, ok}, "stage2": {...}, "y_matches_reference": bool}, every number
    produced by the literal four-step kernel (b200fhe_four_step_protected_host).  `fault` =
    (stage, row, delta) adds delta to one entry of that stage's output, as demo() does (:271-274)."""
    if len(a) != N:
        raise ValueError("len(a) != N")
    plan = [(0, fault[0], fault[1], fault[2])] if fault else []
    y, chk, _ = B.four_step_protected_host(_arr(a, mod).reshape(1, N), mod, g, plan=plan)
    y = [int(v) for v in y[0]]
    l1, r1, l2, r2 = (int(v) for v in chk[0])
    checks = {"stage1": {"lhs": l1, "rhs": r1, "ok": l1 == r1}, "stage2": {"lhs": l2, "rhs": r2, "ok": l2 == r2},
              "y_matches_reference": y == four_step_ntt(a, N, mod, g)}
    return y, checks
