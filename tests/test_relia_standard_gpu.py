"""GPU parity for the baseline scheme of the reference's comparison (b200fhe_relia_standard_host =
non_four_step_ntt_sim.cpp: standard iterative NTT, lazy-reduction bucket checker :135-158, global check :220-235)
  (a) bit for bit against the oracle's restatement, trial by trial (outputs and both detector flags), and
  (b) against the REFERENCE BINARY's printed table (oracle/_ref/non_four_step_ntt_sim --N 64 --q-bits 10): 10^5 trials per
      scenario, 8 scenarios, both detector columns, two-sided binomial 3 sigma."""
import os

import numpy as np
import pytest

from relia_common import SCENARIOS, STD_SIM, draw_faults, run_standard_simulator, within_3_sigma

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,q_bits", [(64, 10), (8, 12), (32, 30), (512, 20), (4096, 30), (1024, 59)])
def test_trials_bit_exact_vs_oracle(pkg, O, N, q_bits):
    q = int(O.lib().orc_find_ntt_prime(N, q_bits))
    g = int(O.lib().orc_find_primitive_root_relia(q))
    lg = N.bit_length() - 1
    total = (N // 2) * lg
    rng = np.random.default_rng(N + q_bits)
    trials = 400 if N <= 512 else 40
    a = rng.integers(0, q, size=(trials, N), dtype=np.uint64)
    w = rng.integers(0, q, size=(trials, N), dtype=np.uint64)
    faults = []
    for t in range(trials):
        if t % 7 == 0:
            continue
        name, kinds = SCENARIOS[t % len(SCENARIOS)]
        f = draw_faults(rng, kinds, 1, 0, total, q.bit_length())
        faults += [(t, int(e["op"]), int(e["kind"]), int(e["payload"])) for e in f]
    out, flags = pkg.binding.relia_standard_host(q, g, a, w, faults)
    per_trial = {}
    for (t, op, kind, payload) in faults:
        per_trial.setdefault(t, []).append((op, kind, payload))
    n_det = 0
    for t in range(trials):
        wo, wf = O.relia_standard(q, g, a[t], w[t], per_trial.get(t, []))
        assert int(flags[t]) == wf, (t, int(flags[t]), wf)
        assert np.array_equal(out[t], wo), t
        if t % 7 == 0:
            assert wf == 0
        n_det += wf != 0
    assert n_det > trials // 2


def test_rejects_what_the_simulator_cannot_run(pkg):
    a = np.zeros((1, 48), dtype=np.uint64)
    with pytest.raises(Exception):
        pkg.binding.relia_standard_host(769, 11, a, a)                    # N not a power of two
    a = np.zeros((1, 64), dtype=np.uint64)
    with pytest.raises(Exception):
        pkg.binding.relia_standard_host(771, 11, a, a)                    # not a prime
    with pytest.raises(Exception):
        pkg.binding.relia_standard_host(769, 11, a, a, [(0, 10 ** 6, 1, 1)])   # op past the last product


@pytest.mark.skipif(not os.path.exists(STD_SIM), reason="oracle/_ref/non_four_step_ntt_sim not built")
def test_detection_rates_match_the_reference_simulator(pkg, O):
    N, q_bits, trials = 64, 10, 100000
    q, table = run_standard_simulator(N, q_bits, trials)
    g = int(O.lib().orc_find_primitive_root_relia(q))
    assert q == 769
    total = (N // 2) * 6
    rng = np.random.default_rng(20251106)
    report = []
    for name, kinds in SCENARIOS:
        a = rng.integers(0, q, size=(trials, N), dtype=np.uint64)
        w = rng.integers(0, q, size=(trials, N), dtype=np.uint64)
        f = draw_faults(rng, kinds, trials, 0, total, q.bit_length())
        _, flags = pkg.binding.relia_standard_host(q, g, a, w, f, want_out=False)
        got = dict(det=int((flags != 0).sum()), lazy=int((flags & 1).sum()), inter=int(((flags >> 1) & 1).sum()))
        ref = table[name]
        report.append((name, got, ref))
        for key in ("det", "lazy", "inter"):
            assert within_3_sigma(got[key], trials, ref[key], ref["trials"]), (name, key, got, ref)
    # the laws behind the numbers: a single bit flip always changes its bucket's sum (lazy catches every SBF / DBF), and the
    # global weighted check misses a propagated single fault with probability ~ 1/Q
    for name, got, ref in report:
        if name in ("SBF", "DBF"):
            assert got["lazy"] == trials
            miss = 1.0 - got["inter"] / trials
            assert 0.4 / q < miss < 2.0 / q, (name, miss)
    out = os.environ.get("B200_RELIA_STD_REPORT")
    if out:
        with open(out, "w") as fh:
            fh.write("# GPU (b200fhe_relia_standard_host) vs reference binary oracle/_ref/non_four_step_ntt_sim, N=64, Q=769, %d trials per scenario\n" % trials)
            fh.write("# scenario | gpu det lazy inter | sim det lazy inter\n")
            for name, got, ref in report:
                fh.write("%-8s | %6d %6d %6d | %6d %6d %6d\n" % (name, got["det"], got["lazy"], got["inter"], ref["det"], ref["lazy"], ref["inter"]))
