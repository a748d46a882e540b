"""CPU tier of SURVEY 8(f) row f4: the oracle's restatement of the Barrett signature / residue fold checks
(oracle.barrett_fold) against first principles and against the tables the REFERENCE binaries printed
(tests/golden/fold_ref_tables.json: rfhe_framewk/src/barrett_final.cpp and barrett_test/outer_check.cpp)."""
import numpy as np

from fold_common import Q37, draw, golden, missed_signature, missed_verdict, within_3_sigma


def test_barrett_params_and_residue_fold(O):
    K, mu = O.barrett_params(Q37)
    assert K == 37 and mu == (1 << 74) // Q37                       # make_barrett_ctx, header line of the reference run
    assert golden()["barrett_final"]["header"] == "q=%d (bits=37), N=64, K=37" % Q37
    rng = np.random.default_rng(3)
    for k in (2, 3, 4, 7, 8, 12, 16, 23, 26):
        M = (1 << k) + 1
        for _ in range(200):
            lo, hi = int(rng.integers(0, 2 ** 63)) * 2 + int(rng.integers(0, 2)), int(rng.integers(0, 2 ** 63))
            assert O.lib().orc_fold_residue(lo, hi, k) == ((hi << 64) | lo) % M
        assert O.lib().orc_fold_residue(0, 0, k) == 0 and O.lib().orc_fold_residue(M, 0, k) == 0


def test_product_is_the_references_barrett(O):
    rng = np.random.default_rng(4)
    for q in (65521, 4294967291, Q37):                               # true modular product while a b < 2^64
        a = rng.integers(0, q, size=(3, 100), dtype=np.uint64)
        b = rng.integers(0, q, size=(3, 100), dtype=np.uint64)
        c, s1, s2, v = O.barrett_fold(q, a, b, fold=7, k=8)
        t = (a.astype(object) * b.astype(object)) % (1 << 64)        # barrett_final.cpp:239 u64 ti = a * b
        assert (c.astype(object) == t % q).all() and not v.any()
        K, _ = O.barrett_params(q)
        want = np.zeros(3 * 15, dtype=np.uint64)
        for vec in range(3):
            for i in range(100):
                want[vec * 15 + i // 7] ^= np.uint64(int(c[vec, i]) & ((1 << K) - 1))
        assert (s2 == want).all()


def test_signature_miss_rates_match_barrett_final(O):
    g = golden()["barrett_final"]
    vec_len, trials = g["args"]["vec_len"], 20000
    rng = np.random.default_rng(11)
    a = rng.integers(0, Q37, size=(trials, vec_len), dtype=np.uint64)
    b = rng.integers(0, Q37, size=(trials, vec_len), dtype=np.uint64)
    for mode in ("SBF", "DBF", "SBF+SBF", "SBF+DBF", "MOF1"):
        f = draw(rng, mode, trials, vec_len, 37)
        fl = [(int(e["index"]), int(e["product"]), int(e["kind"]), int(e["payload"])) for e in f]
        for fold in (1, 8, 64):
            _, c1, c2, _ = O.barrett_fold(Q37, a, b, fold)
            _, f1, f2, _ = O.barrett_fold(Q37, a, b, fold, faults=fl)
            ref = g["%s/instra/%d" % (mode, fold)]
            assert within_3_sigma(missed_signature(c1, f1, trials), trials, ref["missed"], ref["trials"]), (mode, fold)
            assert missed_signature(c2, f2, trials) == trials          # "sum": blind without propagation (miss rate 1.0)
            assert g["%s/sum/%d" % (mode, fold)]["missed"] == g["%s/sum/%d" % (mode, fold)]["trials"]


def test_residue_miss_rates_match_outer_check(O):
    g = golden()["outer_check"]
    vec_len, trials = g["args"]["n"], 30000
    rng = np.random.default_rng(12)
    a = rng.integers(0, Q37, size=(trials, vec_len), dtype=np.uint64)
    b = rng.integers(0, Q37, size=(trials, vec_len), dtype=np.uint64)
    for mode in ("SBF", "DBF", "SBF+SBF", "SBF+DBF", "MOF1", "MOF2"):
        f = draw(rng, mode, trials, vec_len, 37, product=1, mof_both=False)
        fl = [(int(e["index"]), int(e["product"]), int(e["kind"]), int(e["payload"])) for e in f]
        for k in (4, 8, 12, 16):
            _, _, _, v = O.barrett_fold(Q37, a, b, fold=vec_len, k=k, faults=fl)
            for strat, bit in (("intra", 0), ("inter", 1)):
                ref = g["%s/%s/%d" % (mode, strat, k)]
                assert within_3_sigma(missed_verdict(v, trials, bit), trials, ref["missed"], ref["trials"]), (mode, strat, k)
