"""GPU parity of the checksum-carrying base conversion (spec rfhe_framewk/src/baseConv.py:42-98
bConv_ecc): outputs, verdicts and both sides of every check bit-exact against the oracle, fault-free and with
faults injected in each of the three places the checks cover."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _plan(pkg, log_n, bits, m, K):
    N = 1 << log_n
    moduli = pkg.primes_create(N, [bits] * (m + K))
    ctx = pkg.Context(log_n, moduli)
    return N, moduli, pkg.BaseConv(ctx, list(range(m)), list(range(m, m + K)))


@pytest.mark.parametrize("log_n,bits,m,K", [(5, 30, 2, 3), (8, 50, 3, 5), (10, 60, 16, 9), (9, 60, 7, 4), (11, 60, 4, 13)])
def test_fault_free_matches_oracle_and_unprotected_path(pkg, O, log_n, bits, m, K):
    N, moduli, bc = _plan(pkg, log_n, bits, m, K)
    x = O.uniform_limbs(moduli[:m], N, seed=log_n + m)
    out, verdict, sums = bc.modup_protected_host(x)
    w_out, w_verdict, w_sums = O.bconv_hps_protected(x, moduli[:m], moduli[m:])
    assert np.array_equal(out, w_out) and np.array_equal(out, bc.modup_host(x))
    assert not verdict.any() and np.array_equal(verdict, w_verdict)
    assert np.array_equal(sums, w_sums)
    assert np.array_equal(sums[:, 0], sums[:, 1])


@pytest.mark.parametrize("log_n,bits,m,K", [(12, 60, 16, 9), (13, 50, 3, 5), (12, 60, 7, 50), (12, 60, 1, 17), (14, 60, 16, 48)])
def test_tcgen05_contraction_carries_the_sums(pkg, O, log_n, bits, m, K):
    """N >= 4096: the checksum-carrying flavour of the tcgen05 contraction (k_bconv_contract_tc5<true>), incl. ragged chunks (K not a
    multiple of 16), more than 48 targets (second launch: input sums taken once) and a single source limb."""
    N, moduli, bc = _plan(pkg, log_n, bits, m, K)
    x = O.uniform_limbs(moduli[:m], N, seed=log_n + m + K)
    out, verdict, sums = bc.modup_protected_host(x)
    w_out, w_verdict, w_sums = O.bconv_hps_protected(x, moduli[:m], moduli[m:])
    assert np.array_equal(out, w_out) and np.array_equal(out, bc.modup_host(x))
    assert not verdict.any() and np.array_equal(sums, w_sums)
    # faults: a contraction output in the first and in the last target, a y word in memory, a scale product
    inject = [(2, 0, 5, 1 << 11), (2, K - 1, N - 1, 1 << 58), (3, m - 1, 4097 % N, 1 << 30), (1, 0, 77, 1 << 3)]
    out, verdict, sums = bc.modup_protected_host(x, inject=inject)
    w_out, w_verdict, w_sums = O.bconv_hps_protected(x, moduli[:m], moduli[m:], inject=inject)
    assert np.array_equal(out, w_out) and np.array_equal(verdict, w_verdict) and np.array_equal(sums, w_sums)
    assert verdict[m] & 1 and verdict[m + K - 1] & 1 and verdict[m - 1] & 2 and verdict[0] & 1


@pytest.mark.parametrize("inject", [
    [(1, 2, 17, 1 << 40)],                     # scale product of source limb 2
    [(2, 3, 100, 1)],                          # lowest bit of one contraction output of target 3
    [(3, 0, 255, 1 << 59)],                    # y word flipped in memory between the kernels
    [(1, 0, 0, 1), (2, 0, 0, 1 << 33), (3, 5, 9, 1 << 7)],
    [(2, 8, 511, 1 << 20), (2, 8, 510, 1 << 21)],      # two faults in the same target limb (different bits: cannot cancel)
])
def test_injected_faults_verdict_parity(pkg, O, inject):
    N, moduli, bc = _plan(pkg, 9, 60, 6, 9)
    x = O.uniform_limbs(moduli[:6], N, seed=3)
    out, verdict, sums = bc.modup_protected_host(x, inject=inject)
    w_out, w_verdict, w_sums = O.bconv_hps_protected(x, moduli[:6], moduli[6:], inject=inject)
    assert np.array_equal(out, w_out)
    assert np.array_equal(verdict, w_verdict)
    assert np.array_equal(sums, w_sums)
    assert verdict.any()
    # the verdict names the limb that was hit
    for st, limb, _, _ in inject:
        if st == 1:
            assert verdict[limb] & 1
        if st == 3:
            assert verdict[limb] & 2
        if st == 2:
            assert verdict[6 + limb] & 1


def test_full_size_every_single_bit_fault_is_detected(pkg, O):
    """C5's shape (N=2^17, 16 -> 48 limbs): fault-free verdict clean and equal to the unprotected path; one flipped bit
    anywhere (sampled) trips exactly the check of the limb it hit."""
    N, moduli, bc = _plan(pkg, 17, 60, 16, 48)
    x = O.uniform_limbs(moduli[:16], N, seed=17)
    out, verdict, sums = bc.modup_protected_host(x)
    assert not verdict.any()
    assert np.array_equal(out, bc.modup_host(x))
    rng = np.random.default_rng(4)
    for trial in range(12):
        st = int(rng.integers(1, 4))
        limb = int(rng.integers(0, 48 if st == 2 else 16))
        inj = [(st, limb, int(rng.integers(0, N)), 1 << int(rng.integers(0, 59)))]
        _, v, _ = bc.modup_protected_host(x, inject=inj)
        idx = 16 + limb if st == 2 else limb
        assert v[idx] & (2 if st == 3 else 1), inj
        # a stage-1 / stage-3 fault also corrupts the y the contraction consumes, but consistently: its checks stay green
        others = np.delete(v, idx)
        assert not others.any(), inj


def test_prototype_signature_raises_on_mismatch(pkg, O):
    """prototypes.bConv_ecc mirrors baseConv.py:42-98: returns result[i][k], raises ValueError('ECC mismatch ...')."""
    from fhe_reliability_gpu_b200 import prototypes as P
    N = 64
    moduli = pkg.primes_create(N, [40] * 5)
    p, q = moduli[:2], moduli[2:]
    x = O.uniform_limbs(p, N, seed=8)
    res = P.bConv_ecc([[int(v) for v in row] for row in x], p, q)
    want = O.bconv_hps(x, p, q)
    assert res == [[int(want[k][i]) for k in range(3)] for i in range(N)]
    with pytest.raises(ValueError, match="ECC mismatch"):
        P.bConv_ecc([[int(v) for v in row] for row in x], p, q, inject=[(2, 1, 5, 4)])


def test_large_base_is_rejected(pkg, O):
    N, moduli, bc = _plan(pkg, 10, 60, 20, 3)
    x = O.uniform_limbs(moduli[:20], N, seed=1)
    with pytest.raises(pkg.B200FHEError):
        bc.modup_protected_host(x)
