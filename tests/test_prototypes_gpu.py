"""The reference's Python prototype interfaces served by the GPU (fhe_reliability_gpu_b200/prototypes.py)
against golden vectors produced by the prototypes themselves (tests/golden/make_golden.py imports
rfhe_framewk/src/negaclic_ntt.py and reliability_test/four_step_ntt_prot.py from the reference)."""
import json
import os

import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return json.load(open(os.path.join(GOLD, name)))


@pytest.mark.parametrize("i", range(5))
def test_negacyclic_prototype_signatures_match_golden(pkg, i):
    from fhe_reliability_gpu_b200 import prototypes as P
    c = load("negacyclic_py.json")["cases"][i]
    N, q, psi, a = c["N"], c["q"], c["psi"], c["a"]
    A = P.negacyclic_ntt(a, psi, q)
    assert A == c["A_natural"]                                  # negaclic_ntt.py:86-92, natural order
    assert P.negacyclic_intt(A, psi, q) == [x % q for x in a]   # :102-109
    # cyclic transform of the same module (:38-57): ntt(a psi^i, psi^2) is the negacyclic one
    tw = [a[j] * pow(psi, j, q) % q for j in range(N)]
    assert P.ntt(tw, psi * psi % q, q) == c["A_natural"]
    assert P.intt(c["A_natural"], psi * psi % q, q) == tw
    # ecc_check (:130-149): both checksums come out of the fused protected transform
    s_in, s_out, ok = P.ecc_check(a, psi, q)
    assert ok and s_in == s_out == c["ecc_checksum"]
    assert P.generate_weights(N) == c["ecc_w"]


@pytest.mark.parametrize("i", range(4))
def test_four_step_ntt_matches_prototype(pkg, i):
    from fhe_reliability_gpu_b200 import prototypes as P
    c = load("four_step_prot.json")["cases"][i]
    assert P.four_step_ntt(c["a"], c["N"], c["mod"], c["g"]) == c["y"]      # four_step_ntt_prot.py:71-109
    assert P.ntt_direct(c["a"], c["N"]) == c["y"]                            # defaults MOD, G (:16-17)


def test_composite_modulus_is_rejected(pkg):
    """motivation/ntt.py:35-37 runs its inline loop mod 15728641 = 173 * 90917, which is not a prime
    (so not an NTT modulus: no inverse transform exists); the library refuses it instead of computing
    something.  That KAT stays with the oracle's radix-2 loop test (tests/test_oracle_golden.py)."""
    from fhe_reliability_gpu_b200 import prototypes as P
    with pytest.raises(Exception):
        P.ntt([i % 16 for i in range(1024)], pow(3, (15728641 - 1) // 1024, 15728641), 15728641)


def test_ecc_check_flags_an_injected_fault(pkg):
    from fhe_reliability_gpu_b200 import prototypes as P
    c = load("negacyclic_py.json")["cases"][3]
    s_in, s_out, ok = P.ecc_check(c["a"], c["psi"], c["q"], plan=[(0, 3, 17, 1, 1 << 20)])
    assert not ok and s_in == c["ecc_checksum"] and s_out != s_in


def test_bad_arguments_raise(pkg):
    from fhe_reliability_gpu_b200 import prototypes as P
    with pytest.raises(ValueError):
        P.negacyclic_ntt([1, 2, 3], 19, 97)                      # not a power of two
    with pytest.raises(Exception):
        P.negacyclic_ntt([1] * 16, 5, 97)                        # 5 is not a primitive 32nd root mod 97


def test_literal_four_step_protection_checks_match_prototype(pkg):
    """four_step_with_protection_vector (four_step_ntt_prot.py:196-252): y and both (lhs, rhs) pairs equal the
    prototype's own numbers, including demo(16, seed=42): stage1 527369946, stage2 4580965."""
    from fhe_reliability_gpu_b200 import prototypes as P
    g = load("four_step_prot.json")
    n = 0
    for c in g["cases"]:
        y, checks = P.four_step_with_protection_vector(c["a"], c["N"], c["mod"], c["g"])
        assert y == c["y"] and checks["y_matches_reference"]
        if "checks" in c:
            n += 1
            for st in ("stage1", "stage2"):
                assert checks[st]["lhs"] == c["checks"][st]["lhs"] and checks[st]["rhs"] == c["checks"][st]["rhs"] and checks[st]["ok"]
    assert n >= 1
    d16 = [c for c in g["cases"] if c["N"] == 16 and c["seed"] == 42][0]
    y, checks = P.four_step_with_protection_vector(d16["a"], 16)
    assert (checks["stage1"]["lhs"], checks["stage2"]["lhs"]) == (527369946, 4580965)
    # the demo's fault (:271-279): row 13 of the stage-1 output, +34126397 -> lhs 561496343, rhs unchanged, detected
    f = g["demo_fault"]
    y, checks = P.four_step_with_protection_vector(d16["a"], 16, fault=(1, f["row"], f["delta"]))
    assert checks["stage1"] == {"lhs": 561496343, "rhs": 527369946, "ok": False}
    assert not checks["y_matches_reference"]


def test_literal_four_step_batch_sizes_and_stage2_fault(pkg):
    import numpy as np
    from fhe_reliability_gpu_b200 import binding as B, prototypes as P
    mod, gen = 998244353, 3
    rng = np.random.default_rng(4)
    for N in (4, 64, 1024, 4096):
        a = rng.integers(0, mod, size=(3, N), dtype=np.uint64)
        y, chk, verdict = B.four_step_protected_host(a, mod, gen, plan=[(1, 2, N - 1, 5)])
        assert verdict.tolist() == [0, 2, 0]
        assert (chk[:, 0] == chk[:, 1]).all() and chk[0, 2] == chk[0, 3] and chk[1, 2] != chk[1, 3]
        assert [int(v) for v in y[0]] == P.ntt([int(v) for v in a[0]], pow(gen, (mod - 1) // N, mod), mod)
    with pytest.raises(B.B200FHEError):
        B.four_step_protected_host(np.zeros((1, 32), dtype=np.uint64), mod, gen)      # not a perfect square
