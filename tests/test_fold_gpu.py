"""GPU parity for SURVEY 8(f) row f4: the device form of the Barrett signature / residue fold checks
(b200fhe_barrett_fold_host; rfhe_framewk/src/barrett_final.cpp:120-141,175-266, barrett_test/outer_check.cpp:35-50,222-252)
  (a) bit for bit against the oracle's restatement (products, both XOR signatures, every verdict bit), clean and under
      injected product faults, with and without propagation, for ragged fold / vector sizes, and
  (b) its Monte-Carlo miss rates against the tables the REFERENCE binaries printed (tests/golden/fold_ref_tables.json),
      10^5 trials per cell, two-sided binomial 3 sigma."""
import numpy as np
import pytest

from fold_common import Q37, FAULT_DT, draw, golden, missed_signature, missed_verdict, within_3_sigma

pytestmark = pytest.mark.gpu


def _as_list(f):
    return [(int(e["index"]), int(e["product"]), int(e["kind"]), int(e["payload"])) for e in f]


@pytest.mark.parametrize("q,n_vec,vec_len,fold,k", [(Q37, 5, 64, 8, 8), (Q37, 3, 100, 7, 4), (Q37, 2, 4096, 64, 16), (65521, 4, 33, 33, 12),
                                                     (4294967291, 1, 1000, 1, 26), (576460752303423619, 2, 257, 128, 2)])
def test_device_matches_oracle_bit_for_bit(pkg, O, q, n_vec, vec_len, fold, k):
    rng = np.random.default_rng(vec_len * 31 + fold)
    a = rng.integers(0, q, size=(n_vec, vec_len), dtype=np.uint64)
    b = rng.integers(0, q, size=(n_vec, vec_len), dtype=np.uint64)
    K, mu = pkg.binding.barrett_params(q)
    assert (K, mu) == O.barrett_params(q)
    n = n_vec * vec_len
    faults = []
    for j in range(12):
        idx = int(rng.integers(0, n)) if j % 4 else (faults[-1][0] if faults else 0)     # some elements are hit twice
        faults.append((idx, 1 + j % 2, 1 + (j // 2) % 2, int(rng.integers(1, 2 ** 63))))
    for fl in ([], faults):
        for propagate in (0, 1):
            want = O.barrett_fold(q, a, b, fold, k=k, faults=fl, propagate=propagate)
            got = pkg.binding.barrett_fold_host(q, a, b, fold, k=k, faults=fl, propagate=propagate)
            for w, g_, name in zip(want, got, ("c", "sig_intra", "sig_sum", "verdict")):
                assert (w == g_).all(), (name, propagate, len(fl))
            if not fl:
                assert not got[3].any()
            sig_only = pkg.binding.barrett_fold_host(q, a, b, fold, faults=fl, propagate=propagate)
            assert (sig_only[0] == want[0]).all() and (sig_only[1] == want[1]).all() and (sig_only[2] == want[2]).all()
    if q < (1 << 32):
        assert (O.barrett_fold(q, a, b, fold)[0].astype(object) == (a.astype(object) * b.astype(object)) % q).all()


def test_propagated_fault_reaches_the_sum_signature(pkg, O):
    """With propagation the corrupted product changes c, so the cheap "sum" signature sees it too
    (the reference's no-propagation model keeps it blind: barrett_final.cpp:11)."""
    rng = np.random.default_rng(9)
    a = rng.integers(0, Q37, size=(1, 64), dtype=np.uint64); b = rng.integers(0, Q37, size=(1, 64), dtype=np.uint64)
    clean = pkg.binding.barrett_fold_host(Q37, a, b, 8, k=8)
    hit = pkg.binding.barrett_fold_host(Q37, a, b, 8, k=8, faults=[(19, 2, 1, 1 << 5)], propagate=1)
    assert hit[0][0, 19] != clean[0][0, 19] and (hit[0] != clean[0]).sum() == 1
    assert (hit[2] != clean[2]).sum() == 1 and (hit[1] != clean[1]).sum() == 1
    assert hit[3][19 // 8] & 1 and not np.delete(hit[3], 19 // 8).any()


def test_signature_miss_rates_vs_reference_binary(pkg):
    g = golden()["barrett_final"]
    vec_len, trials = g["args"]["vec_len"], 100000
    rng = np.random.default_rng(21)
    a = rng.integers(0, Q37, size=(trials, vec_len), dtype=np.uint64)
    b = rng.integers(0, Q37, size=(trials, vec_len), dtype=np.uint64)
    for fold in (1, 2, 8, 64):
        clean = pkg.binding.barrett_fold_host(Q37, a, b, fold)
        for mode in ("SBF", "DBF", "SBF+SBF", "SBF+DBF", "MOF1", "MOF2"):
            f = draw(rng, mode, trials, vec_len, 37)
            hit = pkg.binding.barrett_fold_host(Q37, a, b, fold, faults=f)
            ref = g["%s/instra/%d" % (mode, fold)]
            assert within_3_sigma(missed_signature(clean[1], hit[1], trials), trials, ref["missed"], ref["trials"]), (mode, fold)
            assert missed_signature(clean[2], hit[2], trials) == trials
            assert (hit[0] == clean[0]).all()


def test_residue_miss_rates_vs_reference_binary(pkg):
    g = golden()["outer_check"]
    vec_len, trials = g["args"]["n"], 100000
    rng = np.random.default_rng(22)
    a = rng.integers(0, Q37, size=(trials, vec_len), dtype=np.uint64)
    b = rng.integers(0, Q37, size=(trials, vec_len), dtype=np.uint64)
    for mode in ("SBF", "DBF", "SBF+SBF", "SBF+DBF", "MOF1", "MOF2"):
        f = draw(rng, mode, trials, vec_len, 37, product=1, mof_both=False)
        for k in (4, 8, 12, 16):
            v = pkg.binding.barrett_fold_host(Q37, a, b, vec_len, k=k, faults=f)[3]
            for strat, bit in (("intra", 0), ("inter", 1)):
                ref = g["%s/%s/%d" % (mode, strat, k)]
                assert within_3_sigma(missed_verdict(v, trials, bit), trials, ref["missed"], ref["trials"]), (mode, strat, k)
