"""ctypes/numpy face of the CPU oracle (oracle/ntt_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product package never
imports this module.  Parity status: see the header of ntt_oracle.c
("parity unpinned" at bit level against SEAL/Phantom, whose sources are absent).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
U64P = C.POINTER(C.c_uint64)


def _host_tag():
    """Identifies the host CPU (model + ISA flags): a -march=native object must never run on another machine."""
    import hashlib
    try:
        txt = open("/proc/cpuinfo").read()
        key = "".join(l for l in txt.splitlines() if l.startswith(("model name", "flags")))[:20000]
    except OSError:
        key = "unknown"
    return hashlib.sha1(key.encode()).hexdigest()[:10]


def build(force=False, native=False):
    """Portable build (x86-64-v3, `make`), or with native=True a -march=native build for THIS host's cores (the timed CPU arm
    of bench.py).  The native object carries the host's CPU tag in its name; falls back to the portable one."""
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "ntt_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    if native:
        nso = os.path.join(_HERE, "liboracle_native_%s.so" % _host_tag())
        try:
            if force or not os.path.exists(nso) or os.path.getmtime(nso) < os.path.getmtime(src):
                subprocess.check_call(["/usr/bin/gcc", "-O3", "-march=native", "-fopenmp", "-fPIC", "-shared", "-o", nso, src],
                                      stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            return nso
        except (OSError, subprocess.CalledProcessError):
            pass
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build(native=os.environ.get("ORACLE_NATIVE") == "1"))
        u64, i32 = C.c_uint64, C.c_int
        for name, res, args in [
            ("orc_mulmod", u64, [u64, u64, u64]),
            ("orc_powmod", u64, [u64, u64, u64]),
            ("orc_invmod", u64, [u64, u64]),
            ("orc_shoup", u64, [u64, u64]),
            ("orc_barrett_ratio", None, [u64, U64P]),
            ("orc_barrett_reduce128", u64, [u64, u64, u64, U64P]),
            ("orc_is_prime", i32, [u64]),
            ("orc_primes_create", i32, [u64, C.POINTER(i32), i32, U64P]),
            ("orc_find_ntt_prime", u64, [u64, i32]),
            ("orc_find_primitive_root_relia", u64, [u64]),
            ("orc_min_primitive_root", u64, [u64, u64]),
            ("orc_bitrev", u64, [u64, i32]),
            ("orc_ntt_tables", None, [i32, u64, u64, U64P, U64P, U64P, U64P]),
            ("orc_ntt_forward_def", None, [U64P, U64P, i32, u64, u64]),
            ("orc_ntt_forward_simple", None, [U64P, i32, u64, U64P]),
            ("orc_ntt_inverse_simple", None, [U64P, i32, u64, U64P]),
            ("orc_ntt_forward_harvey", None, [U64P, i32, u64, U64P, U64P]),
            ("orc_ntt_inverse_harvey", None, [U64P, i32, u64, U64P, U64P]),
            ("orc_ntt_forward_batch", None, [U64P, i32, i32, U64P, U64P, U64P]),
            ("orc_ntt_inverse_batch", None, [U64P, i32, i32, U64P, U64P, U64P]),
            ("orc_ntt_forward_polys", None, [U64P, i32, i32, i32, U64P, U64P, U64P]),
            ("orc_ntt_inverse_polys", None, [U64P, i32, i32, i32, U64P, U64P, U64P]),
            ("orc_relia_four_step", i32, [u64, u64, i32, i32, U64P, U64P, U64P, C.c_void_p, i32]),
            ("orc_relia_standard", i32, [u64, u64, i32, U64P, U64P, U64P, C.c_void_p, i32]),
            ("orc_set_threads", None, [i32]),
            ("orc_get_threads", i32, []),
            ("orc_abft_weights", None, [i32, U64P]),
            ("orc_abft_output_weights", None, [i32, u64, U64P, U64P, U64P]),
            ("orc_ntt_forward_protected", i32, [U64P, i32, u64, U64P, U64P, U64P, C.c_void_p, i32, U64P, U64P]),
            ("orc_negacyclic_mul_naive", None, [U64P, U64P, U64P, u64, u64]),
            ("orc_pmult_accum", None, [U64P, U64P, U64P, i32, i32, i32, u64, U64P]),
            ("orc_bconv_hps", None, [U64P, U64P, i32, i32, u64, U64P, U64P]),
            ("orc_bconv_proto", None, [U64P, U64P, i32, i32, u64, U64P, U64P]),
            ("orc_root_of_unity", u64, [u64, u64, u64]),
            ("orc_cyclic_ntt_direct", None, [U64P, U64P, u64, u64, u64]),
            ("orc_cyclic_ntt_radix2", None, [U64P, u64, u64, u64]),
            ("orc_four_step_cyclic", i32, [U64P, U64P, u64, u64, u64, U64P, U64P]),
            ("orc_protection_check_blockdiag", i32, [U64P, U64P, u64, u64, u64, u64, U64P, U64P]),
            ("orc_barrett_params", i32, [u64, C.POINTER(i32), U64P]),
            ("orc_fold_residue", C.c_uint32, [u64, u64, i32]),
            ("orc_barrett_fold", None, [u64, U64P, U64P, U64P, u64, u64, i32, i32, i32, C.c_void_p, i32, U64P, U64P, C.c_void_p]),
        ]:
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        _LIB = L
    return _LIB


def _p(a):
    assert a.dtype == np.uint64 and a.flags.c_contiguous
    return a.ctypes.data_as(U64P)


def primes_create(N, bits):
    bits = list(bits)
    arr = (C.c_int * len(bits))(*bits)
    out = np.zeros(len(bits), dtype=np.uint64)
    rc = lib().orc_primes_create(N, arr, len(bits), _p(out))
    if rc:
        raise ValueError("primes_create failed rc=%d" % rc)
    return [int(x) for x in out]


def min_root(N, q):
    return int(lib().orc_min_primitive_root(2 * N, q))


class Tables:
    """Per-modulus tables in the layout ntt_test.cu:60-68 uploads."""

    def __init__(self, logn, q, psi=None):
        self.logn, self.q = logn, int(q)
        self.N = 1 << logn
        self.psi = int(psi) if psi else min_root(self.N, self.q)
        self.rp = np.zeros(self.N, dtype=np.uint64)
        self.rps = np.zeros(self.N, dtype=np.uint64)
        self.irp = np.zeros(self.N, dtype=np.uint64)
        self.irps = np.zeros(self.N, dtype=np.uint64)
        lib().orc_ntt_tables(logn, self.q, self.psi, _p(self.rp), _p(self.rps), _p(self.irp), _p(self.irps))


def ntt_forward(a, tab, kind="harvey"):
    a = np.ascontiguousarray(a, dtype=np.uint64).copy()
    if kind == "harvey":
        lib().orc_ntt_forward_harvey(_p(a), tab.logn, tab.q, _p(tab.rp), _p(tab.rps))
    elif kind == "simple":
        lib().orc_ntt_forward_simple(_p(a), tab.logn, tab.q, _p(tab.rp))
    else:
        out = np.zeros_like(a)
        lib().orc_ntt_forward_def(_p(a), _p(out), tab.logn, tab.q, tab.psi)
        a = out
    return a


def ntt_inverse(a, tab, kind="harvey"):
    a = np.ascontiguousarray(a, dtype=np.uint64).copy()
    if kind == "harvey":
        lib().orc_ntt_inverse_harvey(_p(a), tab.logn, tab.q, _p(tab.irp), _p(tab.irps))
    else:
        lib().orc_ntt_inverse_simple(_p(a), tab.logn, tab.q, _p(tab.irp))
    return a


class BatchTables:
    """The per-modulus tables of a limb range stacked ONCE ([L][N] each), for the batched drivers."""

    def __init__(self, tabs):
        self.tabs = list(tabs)
        self.logn, self.L = self.tabs[0].logn, len(self.tabs)
        self.q = np.array([t.q for t in self.tabs], dtype=np.uint64)
        self.rp = np.stack([t.rp for t in self.tabs]); self.rps = np.stack([t.rps for t in self.tabs])
        self.irp = np.stack([t.irp for t in self.tabs]); self.irps = np.stack([t.irps for t in self.tabs])


def _batch(tabs):
    return tabs if isinstance(tabs, BatchTables) else BatchTables(tabs)


def ntt_forward_batch(a, tabs, inplace=False):
    """a: [L][N] or [P][L][N]; limb l of every polynomial uses tabs[l] (a list of Tables, or a BatchTables built once)."""
    a = np.ascontiguousarray(a, dtype=np.uint64)
    if not inplace:
        a = a.copy()
    b = _batch(tabs)
    assert a.shape[-2] == b.L
    lib().orc_ntt_forward_polys(_p(a), b.logn, b.L, a.size >> b.logn, _p(b.q), _p(b.rp), _p(b.rps))
    return a


def ntt_inverse_batch(a, tabs, inplace=False):
    a = np.ascontiguousarray(a, dtype=np.uint64)
    if not inplace:
        a = a.copy()
    b = _batch(tabs)
    assert a.shape[-2] == b.L
    lib().orc_ntt_inverse_polys(_p(a), b.logn, b.L, a.size >> b.logn, _p(b.q), _p(b.irp), _p(b.irps))
    return a


def set_threads(n=None):
    """Pin the OpenMP thread count of the batched drivers (default: every host core) and return what is in force."""
    lib().orc_set_threads(int(n) if n else (os.cpu_count() or 1))
    return int(lib().orc_get_threads())


class OrcInject(C.Structure):
    _fields_ = [("stage", C.c_uint32), ("kind", C.c_uint32), ("butterfly", C.c_uint64), ("payload", C.c_uint64)]


def abft_weights(tab):
    """(w, omega) for one modulus: generate_weights(N) and its transformed weights (SEAL output order)."""
    w = np.zeros(tab.N, dtype=np.uint64); om = np.zeros(tab.N, dtype=np.uint64)
    lib().orc_abft_weights(tab.logn, _p(w))
    lib().orc_abft_output_weights(tab.logn, tab.q, _p(tab.irp), _p(w), _p(om))
    return w, om


def ntt_forward_protected(a, tab, plan=(), weights=None):
    """plan: iterable of (stage, butterfly, kind, payload).  Returns (out, detected, lhs, rhs)."""
    a = np.ascontiguousarray(a, dtype=np.uint64).copy()
    w, om = weights if weights is not None else abft_weights(tab)
    plan = list(plan)
    arr = (OrcInject * max(len(plan), 1))()
    for i, (stage, bf, kind, payload) in enumerate(plan):
        arr[i] = OrcInject(stage, kind, bf, payload)
    lhs, rhs = C.c_uint64(), C.c_uint64()
    det = lib().orc_ntt_forward_protected(_p(a), tab.logn, tab.q, _p(tab.rp), _p(w), _p(om), C.cast(arr, C.c_void_p), len(plan),
                                          C.byref(lhs), C.byref(rhs))
    return a, bool(det), int(lhs.value), int(rhs.value)


def pmult_accum(ct, pt, q):
    """ct [T][P][L][N], pt [T][L][N] -> out [P][L][N]."""
    ct = np.ascontiguousarray(ct, dtype=np.uint64); pt = np.ascontiguousarray(pt, dtype=np.uint64)
    T, P, L, N = ct.shape
    out = np.zeros((P, L, N), dtype=np.uint64)
    qa = np.array(q, dtype=np.uint64)
    lib().orc_pmult_accum(_p(ct), _p(pt), _p(out), T, P, L, N, _p(qa))
    return out


def bconv_hps(x, p, qo):
    """x [m][N] mod p_j -> [K][N] mod qo_k (HPS fast base conversion)."""
    x = np.ascontiguousarray(x, dtype=np.uint64)
    m, N = x.shape
    out = np.zeros((len(qo), N), dtype=np.uint64)
    pa = np.array(p, dtype=np.uint64); qa = np.array(qo, dtype=np.uint64)
    lib().orc_bconv_hps(_p(x), _p(out), m, len(qo), N, _p(pa), _p(qa))
    return out


def bconv_proto(x, p, qo):
    """rfhe_framewk/src/baseConv.py:10-40 bConv: returns result[i][k]."""
    x = np.ascontiguousarray(x, dtype=np.uint64)
    m, N = x.shape
    out = np.zeros((N, len(qo)), dtype=np.uint64)
    pa = np.array(p, dtype=np.uint64); qa = np.array(qo, dtype=np.uint64)
    lib().orc_bconv_proto(_p(x), _p(out), m, len(qo), N, _p(pa), _p(qa))
    return out


def cyclic_ntt_direct(a, N, mod, g):
    a = np.ascontiguousarray(a, dtype=np.uint64); y = np.zeros(N, dtype=np.uint64)
    lib().orc_cyclic_ntt_direct(_p(a), _p(y), N, mod, g)
    return y


def cyclic_ntt_radix2(a, mod, root):
    """motivation/ntt.py:8-32 ntt(a, mod, root)."""
    a = np.ascontiguousarray(a, dtype=np.uint64).copy()
    lib().orc_cyclic_ntt_radix2(_p(a), len(a), mod, root)
    return a


def four_step_cyclic(a, N, mod, g):
    """four_step_ntt_prot.py:71-109; returns (y, B after step 1, C after twiddle)."""
    a = np.ascontiguousarray(a, dtype=np.uint64)
    y = np.zeros(N, dtype=np.uint64); B = np.zeros(N, dtype=np.uint64); Cc = np.zeros(N, dtype=np.uint64)
    rc = lib().orc_four_step_cyclic(_p(a), _p(y), N, mod, g, _p(B), _p(Cc))
    if rc:
        raise ValueError("N must be a perfect square")
    return y, B, Cc


def splitmix64(x):
    x = (x + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
    z = x
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
    return z ^ (z >> 31)


def uniform_limbs(moduli, N, seed=20250606):
    """[L][N] with limb l i.i.d. uniform on [0,q_l) (numpy Generator, seeded)."""
    rng = np.random.default_rng(seed)
    out = np.empty((len(moduli), N), dtype=np.uint64)
    for l, q in enumerate(moduli):
        out[l] = rng.integers(0, int(q), size=N, dtype=np.uint64)
    return out


# ---------------------------------------------------------------- hybrid key-switch
def keyswitch_modup(c, moduli, n_q, n_p, dnum, tabs):
    """ModUp of every digit (dotprod_test.cu:114,146 relinearize/rotate, SEAL switch_key_inplace; spec of the
    conversion itself rfhe_framewk/src/baseConv.py:10-40 in its HPS form): c [n_q][N] NTT domain ->
    [dnum][n_q+n_p][N] NTT domain, the digit's own limbs copied through.  Composed from the pinned primitives."""
    c = np.ascontiguousarray(c, dtype=np.uint64)
    N = c.shape[1]
    alpha, total = n_q // dnum, n_q + n_p
    x = ntt_inverse_batch(c, tabs[:n_q])
    up = np.zeros((dnum, total, N), dtype=np.uint64)
    for d in range(dnum):
        dig = list(range(d * alpha, (d + 1) * alpha))
        tgt = [g for g in range(total) if g not in dig]
        conv = bconv_hps(x[dig], [moduli[g] for g in dig], [moduli[g] for g in tgt])
        up[d, tgt] = ntt_forward_batch(conv, [tabs[g] for g in tgt])
        up[d, dig] = c[dig]
    return up


def bconv_hps_scaled(x, p, qo, t=1):
    """t [x t^-1]_P (+ e t P) mod qo_k: the HPS conversion with a scale folded in (== t * bconv_hps(x t^-1 mod p_j))."""
    if t == 1:
        return bconv_hps(x, p, qo)
    x = np.ascontiguousarray(x, dtype=np.uint64)
    xs = np.stack([np.array((x[j].astype(object) * pow(int(t), -1, int(pj))) % int(pj), dtype=np.uint64) for j, pj in enumerate(p)])
    out = bconv_hps(xs, p, qo)
    return np.stack([np.array((out[k].astype(object) * int(t)) % int(qk), dtype=np.uint64) for k, qk in enumerate(qo)])


def keyswitch_moddown(acc, moduli, n_q, n_p, tabs, t=1):
    """acc [n_q+n_p][N] NTT domain -> [n_q][N]: (acc - BaseConv_{P->Q}(acc mod P)) * P^-1 mod q_l.  t > 1: BGV flavour,
    the subtracted term is t [acc t^-1]_P (congruent to acc mod P, to 0 mod t)."""
    acc = np.ascontiguousarray(acc, dtype=np.uint64)
    tt = ntt_inverse_batch(acc[n_q:], tabs[n_q:])
    conv = ntt_forward_batch(bconv_hps_scaled(tt, moduli[n_q:], moduli[:n_q], t), tabs[:n_q])
    P = 1
    for p in moduli[n_q:]:
        P *= int(p)
    N = acc.shape[1]
    qa = np.array([int(q) for q in moduli[:n_q]], dtype=np.uint64)[:, None]
    pinv = np.stack([np.full(N, pow(P % int(q), -1, int(q)), dtype=np.uint64) for q in moduli[:n_q]])
    neg = (qa - conv) % qa
    # (acc - conv) * P^-1 = acc * pinv + (q - conv) * pinv, through the C multiply-accumulate
    return pmult_accum(np.stack([acc[:n_q], neg])[:, None], np.stack([pinv, pinv]), moduli[:n_q])[0]


def keyswitch_hybrid(c, evk, moduli, n_q, n_p, dnum, tabs, t=1):
    """c [n_q][N]; evk [dnum][2][n_q+n_p][N]; returns [2][n_q][N] (all NTT domain).  t: plain modulus (BGV) or 1."""
    up = keyswitch_modup(c, moduli, n_q, n_p, dnum, tabs)
    acc = pmult_accum(np.ascontiguousarray(evk, dtype=np.uint64), up, moduli)       # terms = digits, "plaintext" = up_d
    return np.stack([keyswitch_moddown(acc[p], moduli, n_q, n_p, tabs, t) for p in range(2)])


# ---------------------------------------------------------------- checksum-carrying base conversion
def bconv_hps_protected(x, p, qo, inject=()):
    """HPS conversion with the linearity checks of rfhe_framewk/src/baseConv.py:42-98 (bConv_ecc) applied per step:
    scale (sum y_j == inv_j sum x_j mod p_j), contraction (sum out_k == sum_j hat_jk sum y_j mod q_k), and y unchanged
    between the two.  inject: (stage, limb, coeff, xor_mask); stage 1 = scale product, 2 = contraction output,
    3 = y in memory between the steps.  Returns (out [K][N], verdict [m+K], sums [m+K][2]).  Python integers: small N."""
    x = np.ascontiguousarray(x, dtype=np.uint64)
    m, N = x.shape
    K = len(qo)
    p = [int(v) for v in p]; qo = [int(v) for v in qo]
    P = 1
    for v in p:
        P *= v
    xs = [[int(v) for v in row] for row in x]
    inv = [pow((P // p[j]) % p[j], -1, p[j]) for j in range(m)]
    y = [[(xs[j][i] * inv[j]) % p[j] for i in range(N)] for j in range(m)]
    for st, limb, coeff, mask in inject:
        if st == 1:
            y[limb][coeff] ^= int(mask)
    sum_x = [sum(row) for row in xs]
    sum_y_prod = [sum(row) for row in y]
    for st, limb, coeff, mask in inject:
        if st == 3:
            y[limb][coeff] ^= int(mask)
    sum_y = [sum(row) for row in y]
    hat = [[(P // p[j]) % qo[k] for j in range(m)] for k in range(K)]
    out = np.zeros((K, N), dtype=np.uint64)
    for k in range(K):
        col = [sum(y[j][i] * hat[k][j] for j in range(m)) % qo[k] for i in range(N)]
        for st, limb, coeff, mask in inject:
            if st == 2 and limb == k:
                col[coeff] ^= int(mask)
        out[k] = np.array(col, dtype=np.uint64)
    verdict = np.zeros(m + K, dtype=np.uint32)
    sums = np.zeros((m + K, 2), dtype=np.uint64)
    for j in range(m):
        lhs, rhs = sum_y_prod[j] % p[j], (sum_x[j] % p[j]) * inv[j] % p[j]
        verdict[j] = (1 if lhs != rhs else 0) | (2 if sum_y_prod[j] != sum_y[j] else 0)
        sums[j] = (lhs, rhs)
    for k in range(K):
        lhs = sum(int(v) for v in out[k]) % qo[k]
        rhs = sum((sum_y[j] % qo[k]) * hat[k][j] for j in range(m)) % qo[k]
        verdict[m + k] = 1 if lhs != rhs else 0
        sums[m + k] = (lhs, rhs)
    return out, verdict, sums


# ---------------------------------------------------------------- evaluator steps around the key-switch
def galois_coeff(a, g, q):
    """x -> x^g on a coefficient-domain polynomial mod (x^N + 1, q): coefficient i moves to i g mod 2N, negated when it
    wraps past N (SEAL GaloisTool::apply_galois; the permutation under rotate_inplace, dotprod_test.cu:146)."""
    a = [int(v) for v in a]
    N = len(a)
    out = [0] * N
    for i, v in enumerate(a):
        idx = (i * g) % (2 * N)
        if idx >= N:
            out[idx - N] = (-v) % q
        else:
            out[idx] = v % q
    return np.array(out, dtype=np.uint64)


def ct_tensor(a, b, moduli):
    """a, b [2][L][N] NTT form -> [3][L][N] = (a0 b0, a0 b1 + a1 b0, a1 b1) (phantom::multiply, dotprod_test.cu:113)."""
    a = np.ascontiguousarray(a, dtype=np.uint64); b = np.ascontiguousarray(b, dtype=np.uint64)
    d0 = pmult_accum(a[0][None, None], b[0][None], moduli)[0]
    d2 = pmult_accum(a[1][None, None], b[1][None], moduli)[0]
    d1 = pmult_accum(np.stack([a[0], a[1]])[:, None], np.stack([b[1], b[0]]), moduli)[0]
    return np.stack([d0, d1, d2])


class OrcReliaFault(C.Structure):
    _fields_ = [("op", C.c_uint64), ("kind", C.c_uint32), ("pad", C.c_uint32), ("payload", C.c_uint64)]


def relia_four_step(q, g, a, weights, faults=(), fold_bits=4):
    """One trial of relia_ntt_sim.cpp four_step_ntt: a [N], weights [2 n2 + n1], faults (op, kind, payload).
    Returns (out [N], flags): bit 0 batch 1 failed, 1 intra, 2 inter, 3 batch 2."""
    a = np.ascontiguousarray(a, dtype=np.uint64)
    weights = np.ascontiguousarray(weights, dtype=np.uint64)
    faults = list(faults)
    arr = (OrcReliaFault * max(len(faults), 1))()
    for i, (op, kind, payload) in enumerate(faults):
        arr[i] = OrcReliaFault(int(op), int(kind), 0, int(payload))
    out = np.zeros_like(a)
    flags = lib().orc_relia_four_step(q, g, a.size, fold_bits, _p(a), _p(out), _p(weights), C.cast(arr, C.c_void_p), len(faults))
    return out, int(flags)


def relia_standard(q, g, a, weights, faults=()):
    """One trial of non_four_step_ntt_sim.cpp: a [N], weights [N] (the global check's random w), faults (op, kind, payload).
    Returns (out [N], flags): bit 0 the lazy bucket check failed, bit 1 the global check failed."""
    a = np.ascontiguousarray(a, dtype=np.uint64)
    weights = np.ascontiguousarray(weights, dtype=np.uint64)
    faults = list(faults)
    arr = (OrcReliaFault * max(len(faults), 1))()
    for i, (op, kind, payload) in enumerate(faults):
        arr[i] = OrcReliaFault(int(op), int(kind), 0, int(payload))
    out = np.zeros_like(a)
    flags = lib().orc_relia_standard(q, g, a.size, _p(a), _p(out), _p(weights), C.cast(arr, C.c_void_p), len(faults))
    return out, int(flags)


def bconv_exact(x, p, qo):
    """motivation/baseConv.py:67-83 base_conv_fixed restated with Python integers (small cases): x [m][N] mod p_j ->
    [K][N] = (CRT(x) mod P) mod qo_k."""
    x = np.asarray(x, dtype=np.uint64)
    m, N = x.shape
    P = 1
    for pj in p:
        P *= int(pj)
    hat = [P // int(pj) for pj in p]
    inv = [pow(hat[j] % int(p[j]), -1, int(p[j])) for j in range(m)]
    out = np.zeros((len(qo), N), dtype=np.uint64)
    for i in range(N):
        v = sum((int(x[j, i]) % int(p[j])) * hat[j] * inv[j] for j in range(m)) % P
        for k, q in enumerate(qo):
            out[k, i] = v % int(q)
    return out


class _FoldFault(C.Structure):
    _fields_ = [("index", C.c_uint64), ("payload", C.c_uint64), ("product", C.c_uint32), ("kind", C.c_uint32)]


def barrett_params(q):
    """make_barrett_ctx (barrett_final.cpp:74-80): (K, mu)."""
    K, mu = C.c_int(), C.c_uint64()
    lib().orc_barrett_params(q, C.byref(K), C.byref(mu))
    return int(K.value), int(mu.value)


def barrett_fold(q, a, b, fold, k=None, faults=(), propagate=0):
    """Element-wise Barrett products of a, b [n_vec][vec_len] with the group signatures / residue verdicts of f4.
    faults: (index, product, kind, payload).  Returns (c, sig_intra, sig_sum, verdict or None)."""
    a = np.ascontiguousarray(a, dtype=np.uint64); b = np.ascontiguousarray(b, dtype=np.uint64)
    n_vec, vec_len = a.shape
    groups = n_vec * ((vec_len + fold - 1) // fold)
    c = np.zeros_like(a)
    s1, s2 = np.zeros(groups, dtype=np.uint64), np.zeros(groups, dtype=np.uint64)
    v = np.zeros(groups, dtype=np.uint32) if k else None
    faults = list(faults)
    arr = (_FoldFault * max(len(faults), 1))()
    for i, (idx, prod, kind, payload) in enumerate(faults):
        arr[i] = _FoldFault(int(idx), int(payload), int(prod), int(kind))
    lib().orc_barrett_fold(q, _p(a), _p(b), _p(c), n_vec, vec_len, fold, k or 0, propagate, C.cast(arr, C.c_void_p), len(faults),
                           _p(s1), _p(s2), v.ctypes.data_as(C.c_void_p) if k else None)
    return c, s1, s2, v
