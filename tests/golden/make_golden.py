#!/usr/bin/env python
"""Generates tests/golden/*.json by IMPORTING the reference's own Python
prototypes from /root/reference (authoring container only; the GPU box has no
/root/reference, so the vectors are committed).  Deterministic: every input is
drawn from random.Random(seed) with the seed recorded in the fixture.

Sources (reference paths):
  reliability_test/four_step_ntt_prot.py   four_step_ntt, ntt_direct, protection checks, demo(16, 42)
  rfhe_framewk/src/negaclic_ntt.py         negacyclic_ntt / negacyclic_intt / naive negacyclic product / ecc weights
  rfhe_framewk/src/baseConv.py             bConv, bConv_ecc
  motivation/baseConv.py:67-83             base_conv_fixed (exact CRT conversion)
  motivation/ntt.py:8-32                   inline cyclic radix-2 NTT (restated call; module has plotting side effects)
Run:  python tests/golden/make_golden.py
"""
import importlib.util
import json
import os
import random
import sys

REF = os.environ.get("REF", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))


def load(path, name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, path))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def dump(name, obj):
    with open(os.path.join(HERE, name), "w") as f:
        json.dump(obj, f, separators=(",", ":"))
    print("wrote", name, os.path.getsize(os.path.join(HERE, name)), "bytes")


def min_psi(N, q):
    """smallest primitive 2N-th root (SEAL try_minimal_primitive_root), brute force for small cases"""
    g = 2
    while True:
        r = pow(g, (q - 1) // (2 * N), q)
        if pow(r, N, q) == q - 1:
            break
        g += 1
    best, cur, gen = r, r, pow(r, 2, q)
    for _ in range(N):
        cur = cur * gen % q
        best = min(best, cur)
    return best


def four_step():
    fs = load("reliability_test/four_step_ntt_prot.py", "fs_prot")
    cases = []
    for N, seed in [(16, 42), (16, 1), (64, 7), (256, 3)]:
        rng = random.Random(seed)
        a = [rng.randrange(fs.MOD) for _ in range(N)]
        y4 = fs.four_step_ntt(a, N)
        case = {"N": N, "seed": seed, "mod": fs.MOD, "g": fs.G, "a": a, "y": y4}
        if N <= 64:
            yd = fs.ntt_direct(a, N)
            assert yd == y4
            y, checks = fs.four_step_with_protection_vector(a, N)
            assert y == y4
            case["checks"] = checks
        cases.append(case)
    # demo(16, seed=42) fault line: same RNG stream as the reference demo
    rng = random.Random(42)
    a = [rng.randrange(fs.MOD) for _ in range(16)]
    i = rng.randrange(16)
    delta = rng.randrange(1, fs.MOD)
    dump("four_step_prot.json", {"cases": cases, "demo_fault": {"N": 16, "seed": 42, "row": i, "delta": delta}})


def negacyclic():
    ng = load("rfhe_framewk/src/negaclic_ntt.py", "negaclic_ntt")
    out = []
    # (N, q): q = 1 mod 2N.  The 60-bit / 50-bit entries are the SEAL-rule primes the harness would use.
    for N, q, seed in [(16, 97, 1), (64, 769, 2), (64, 1152921504606844417, 3), (256, 1125899906826241, 4),
                       (1024, 1152921504606830593, 5)]:
        if (q - 1) % (2 * N):
            continue
        psi = min_psi(N, q)
        rng = random.Random(seed)
        a = [rng.randrange(q) for _ in range(N)]
        b = [rng.randrange(q) for _ in range(N)]
        A = ng.negacyclic_ntt(a, psi, q)              # natural order: A[k] = sum a_i psi^{(2k+1)i}
        assert ng.negacyclic_intt(A, psi, q) == a
        case = {"N": N, "q": q, "psi": psi, "seed": seed, "a": a, "A_natural": A}
        if N <= 256:
            case["b"] = b
            case["ab_negacyclic"] = ng.poly_mul_naive_negacyclic(a, b, q)
            assert ng.poly_mul_negacyclic_ntt(a, b, psi, q) == case["ab_negacyclic"]
        if int(N ** 0.5) ** 2 == N:
            w = ng.generate_weights(N)
            psi_inv = ng.modinv(psi, q)
            w_pre = [(w[i] * pow(psi_inv, i, q)) % q for i in range(N)]
            w_hat = ng.intt(w_pre, pow(psi, 2, q), q)
            case["ecc_w"] = w
            case["ecc_w_hat"] = w_hat
            case["ecc_checksum"] = sum(w[i] * a[i] for i in range(N)) % q
            assert case["ecc_checksum"] == sum(w_hat[i] * A[i] for i in range(N)) % q
        out.append(case)
    dump("negacyclic_py.json", {"cases": out})


def baseconv():
    bc = load("rfhe_framewk/src/baseConv.py", "baseConv_ref")
    out = []
    # bConv on general output moduli (baseConv.py:10-40)
    for m, K, N, bits, seed in [(3, 2, 8, 12, 1), (4, 3, 16, 20, 2), (3, 4, 8, 30, 3)]:
        random.seed(seed)
        primes = bc.generate_crt_primes(m + K, bits)
        p, qo = primes[:m], primes[m:]
        res = bc.generate_residues(p, N)
        r1 = bc.bConv([list(r) for r in res], p, qo)
        out.append({"m": m, "K": K, "N": N, "seed": seed, "p": p, "q_out": qo, "residues": res, "bConv": r1})
    # bConv_ecc in the reference's own configuration (baseConv.py:196-214): output moduli are products of
    # groups of input moduli, the only setting in which its ECC identity holds (it raises otherwise).
    for m, K, N, bits, seed in [(4, 2, 16, 10, 4), (4, 2, 32, 12, 5), (6, 3, 16, 10, 6)]:
        random.seed(seed)
        p = bc.generate_crt_primes(m, bits)
        qo = bc.group_multiply(p, m // K)
        res = bc.generate_residues(p, N)
        r1 = bc.bConv([list(r) for r in res], p, qo)
        r2 = bc.bConv_ecc([list(r) for r in res], p, qo)     # N+1 rows; returning at all = ECC check passed
        assert [list(x) for x in r2[:N]] == [list(x) for x in r1]
        out.append({"m": m, "K": K, "N": N, "seed": seed, "p": p, "q_out": qo, "residues": res, "bConv": r1,
                    "bConv_ecc_last": r2[N]})
    dump("baseconv_py.json", {"cases": out})


def motivation_baseconv():
    # motivation/baseConv.py:67-83 base_conv_fixed (the exact-CRT conversion), imported with matplotlib stubbed out
    # (the module imports pyplot at the top for its plotting demo; the package is not installed here)
    import types
    stub = types.ModuleType("matplotlib"); stub.pyplot = types.ModuleType("matplotlib.pyplot")
    sys.modules.setdefault("matplotlib", stub); sys.modules.setdefault("matplotlib.pyplot", stub.pyplot)
    mb = load("motivation/baseConv.py", "motivation_baseConv")
    out = []
    for m, K, N, bits, seed in [(3, 3, 20, 30, 1), (4, 2, 16, 20, 2), (2, 5, 8, 40, 3), (6, 3, 12, 28, 4)]:
        random.seed(seed)
        p = mb.generate_crt_primes(m, bits)
        qo = mb.generate_crt_primes(K, bits)
        P = 1
        for v in p:
            P *= v
        values = [random.randint(0, P - 1) for _ in range(N)]
        values[0], values[1], values[2] = 0, 1, P - 1           # the cases a floating-point estimate of e gets wrong
        res = mb.values_to_rns(values, p)
        conv = mb.base_conv_fixed(res, p, qo)
        assert conv == mb.values_to_rns(values, qo)
        out.append({"m": m, "K": K, "N": N, "seed": seed, "p": p, "q_out": qo, "residues": res, "base_conv_fixed": conv})
    dump("motivation_baseconv.json", {"cases": out})


def motivation_ntt():
    # motivation/ntt.py:8-32 executed on its own deterministic input (:35-41); the module itself
    # imports matplotlib with a font that is absent, so the function source is exec'd in isolation.
    src = open(os.path.join(REF, "motivation/ntt.py")).read().split("# -------------------")[0]
    src = "\n".join(l for l in src.splitlines() if "plt" not in l and "matplotlib" not in l)
    ns = {}
    exec(src, ns)
    n, mod, root = 1024, 15728641, 3
    a = [(i % 16) for i in range(n)]
    A = ns["ntt"](list(a), mod, root)
    dump("motivation_ntt.json", {"n": n, "mod": mod, "root": root, "A": A})


if __name__ == "__main__":
    if not os.path.isdir(REF):
        sys.exit("reference tree absent: fixtures are committed, nothing to do")
    four_step()
    negacyclic()
    baseconv()
    motivation_baseconv()
    motivation_ntt()
