"""CPU tier: the oracle's restatement of non_four_step_ntt_sim.cpp's trial (oracle.relia_standard: standard NTT + lazy bucket
check + global check) against the REFERENCE BINARY oracle/_ref/non_four_step_ntt_sim: per fault scenario the counts of both
detectors agree within binomial 3 sigma (different RNG streams, same law)."""
import os

import numpy as np
import pytest

from relia_common import SCENARIOS, STD_SIM, draw_faults, run_standard_simulator, within_3_sigma


@pytest.mark.skipif(not os.path.exists(STD_SIM), reason="oracle/_ref/non_four_step_ntt_sim not built (reference tree absent at build time)")
def test_oracle_trial_matches_reference_simulator_rates(O):
    N, q_bits, trials = 64, 10, 6000
    q, table = run_standard_simulator(N, q_bits, 60000)
    assert q == 769 == O.lib().orc_find_ntt_prime(N, q_bits)
    g = int(O.lib().orc_find_primitive_root_relia(q))
    total = (N // 2) * 6
    rng = np.random.default_rng(77)
    for name, kinds in SCENARIOS:
        nf = len(kinds)
        f = draw_faults(rng, kinds, trials, 0, total, q.bit_length())
        a = rng.integers(0, q, size=(trials, N), dtype=np.uint64)
        w = rng.integers(0, q, size=(trials, N), dtype=np.uint64)
        cnt = dict(det=0, lazy=0, inter=0)
        for t in range(trials):
            fl = [(int(e["op"]), int(e["kind"]), int(e["payload"])) for e in f[t * nf:(t + 1) * nf]]
            _, flags = O.relia_standard(q, g, a[t], w[t], fl)
            cnt["det"] += flags != 0; cnt["lazy"] += flags & 1; cnt["inter"] += (flags >> 1) & 1
        ref = table[name]
        for key in ("det", "lazy", "inter"):
            assert within_3_sigma(cnt[key], trials, ref[key], ref["trials"]), (name, key, cnt, ref)


@pytest.mark.parametrize("N,q", [(64, 769), (8, 17), (256, 7681)])
def test_fault_free_trial_is_the_cyclic_dft_and_raises_nothing(O, N, q):
    g = int(O.lib().orc_find_primitive_root_relia(q))
    rng = np.random.default_rng(N)
    a = rng.integers(0, q, size=N, dtype=np.uint64)
    w = rng.integers(0, q, size=N, dtype=np.uint64)
    out, flags = O.relia_standard(q, g, a, w)
    assert flags == 0
    wN = pow(g, (q - 1) // N, q)
    want = [sum(int(a[j]) * pow(wN, j * k, q) for j in range(N)) % q for k in range(N)]
    assert [int(x) for x in out] == want


def test_a_reduction_fault_trips_its_bucket_and_only_cancelling_pairs_escape(O):
    q, N = 769, 64
    g = int(O.lib().orc_find_primitive_root_relia(q))
    rng = np.random.default_rng(5)
    a = rng.integers(0, q, size=N, dtype=np.uint64)
    w = rng.integers(0, q, size=N, dtype=np.uint64)
    clean, _ = O.relia_standard(q, g, a, w)
    # a single flipped bit always changes the stored product, so the lazy check (exact sums mod q) always fires
    for op in (0, 31, 32, 100, 191):
        out, flags = O.relia_standard(q, g, a, w, [(op, 1, 1 << (op % 10))])
        assert flags & 1, op
        assert not np.array_equal(out, clean)
    # replacing a product by itself (MLF drawing the true value) is no fault at all: nothing fires
    out, flags = O.relia_standard(q, g, a, w, [(5, 2, int(0))])
    # payload 0: v becomes 0; fires unless the true product was 0
    assert (flags & 1) or np.array_equal(out, clean)
