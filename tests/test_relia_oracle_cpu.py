"""CPU tier: the oracle's restatement of relia_ntt_sim.cpp's trial (oracle.relia_four_step) against the REFERENCE BINARY
oracle/_ref/relia_ntt_sim (built by oracle/Makefile from the reference source where it lies): per fault scenario and per
target stage, the detection counts of every detector agree within binomial 3 sigma (different RNG streams, same law)."""
import os

import numpy as np
import pytest

from relia_common import SCENARIOS, STAGES, SIM, draw_faults, run_simulator, within_3_sigma


def _oracle_cell(O, rng, q, g, N, kinds, lo, hi, trials, fold_bits):
    n1 = int(round(N ** 0.5)); n2 = N // n1
    bits = q.bit_length()
    f = draw_faults(rng, kinds, trials, lo, hi, bits)
    nf = len(kinds)
    a = rng.integers(0, q, size=(trials, N), dtype=np.uint64)
    w = rng.integers(0, q, size=(trials, 2 * n2 + n1), dtype=np.uint64)
    cnt = dict(det=0, b1=0, intra=0, inter=0, b2=0, trials=trials)
    for t in range(trials):
        fl = [(int(e["op"]), int(e["kind"]), int(e["payload"])) for e in f[t * nf:(t + 1) * nf]]
        _, flags = O.relia_four_step(q, g, a[t], w[t], fl, fold_bits)
        cnt["det"] += flags != 0
        cnt["b1"] += flags & 1; cnt["intra"] += (flags >> 1) & 1; cnt["inter"] += (flags >> 2) & 1; cnt["b2"] += (flags >> 3) & 1
    return cnt


@pytest.mark.skipif(not os.path.exists(SIM), reason="oracle/_ref/relia_ntt_sim not built (reference tree absent at build time)")
def test_oracle_trial_matches_reference_simulator_rates(O):
    N, q_bits, fold_bits, trials = 64, 10, 4, 4000
    q, table = run_simulator(N, q_bits, 40000, fold_bits)
    assert q == 769 == O.lib().orc_find_ntt_prime(N, q_bits)          # reliaNTT_test_result.txt:1
    g = int(O.lib().orc_find_primitive_root_relia(q))
    n1 = 8
    s1 = n1 * (8 // 2) * 3; s2 = s1 + N; s3 = s2 + 8 * (n1 // 2) * 3
    ranges = dict(zip(STAGES, [(0, s1), (s1, s2), (s2, s3)]))
    rng = np.random.default_rng(2025)
    for name, kinds in SCENARIOS:
        for stage in STAGES:
            got = _oracle_cell(O, rng, q, g, N, kinds, *ranges[stage], trials, fold_bits)
            ref = table[(name, stage)]
            for key in ("det", "b1", "intra", "inter", "b2"):
                assert within_3_sigma(got[key], trials, ref[key], ref["trials"]), (name, stage, key, got, ref)


def test_fault_free_trial_is_the_cyclic_dft_and_raises_nothing(O):
    q, N = 769, 64
    g = int(O.lib().orc_find_primitive_root_relia(q))
    rng = np.random.default_rng(1)
    a = rng.integers(0, q, size=N, dtype=np.uint64)
    w = rng.integers(0, q, size=24, dtype=np.uint64)
    out, flags = O.relia_four_step(q, g, a, w)
    assert flags == 0
    wN = pow(g, (q - 1) // N, q)
    # four-step index map of the simulator: input A[c n2 + r], output C[r n1 + c']... checked through the DFT definition
    # X[k] = sum_j a_j wN^(jk) for a permutation-free comparison: multiset of outputs of the 2D transform of the 8x8 array
    n1 = n2 = 8
    A = a.reshape(n1, n2).astype(object)
    w2, w1 = pow(wN, n1, q), pow(wN, n2, q)
    B = np.array([[sum(int(A[c][j]) * pow(w2, j * k, q) for j in range(n2)) % q for k in range(n2)] for c in range(n1)], dtype=object)
    B = np.array([[B[c][r] * pow(wN, r * c, q) % q for r in range(n2)] for c in range(n1)], dtype=object)
    C = np.array([[sum(int(B[j][r]) * pow(w1, j * k, q) for j in range(n1)) % q for k in range(n1)] for r in range(n2)], dtype=object)
    assert [int(x) for x in out] == [int(x) for x in C.reshape(-1)]
