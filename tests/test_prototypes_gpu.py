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
