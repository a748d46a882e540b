"""Shared helpers of the f4 (intra-multiplier fold check) tests: the fault scenarios of rfhe_framewk/src/barrett_final.cpp
(:175-232) and rfhe_framewk/src/barrett_test/outer_check.cpp (:113-176) drawn with numpy, and the tables the reference
binaries printed (tests/golden/fold_ref_tables.json, written by tests/golden/make_fold_golden.py)."""
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden", "fold_ref_tables.json")
Q37 = 68719476767                      # next_prime_ge(2^36): the modulus both simulators use (barrett_final.cpp:301)
FAULT_DT = np.dtype([("index", "<u8"), ("payload", "<u8"), ("product", "<u4"), ("kind", "<u4")])


def golden():
    with open(GOLDEN) as f:
        return json.load(f)


def _bits(rng, trials, K, n):
    """n distinct bit positions below K per trial -> OR-ed masks"""
    b1 = rng.integers(0, K, size=trials)
    m = np.left_shift(np.uint64(1), b1.astype(np.uint64))
    if n == 2:
        b2 = rng.integers(0, K - 1, size=trials)
        b2 = np.where(b2 >= b1, b2 + 1, b2)
        m = m | np.left_shift(np.uint64(1), b2.astype(np.uint64))
    return m


def draw(rng, mode, trials, vec_len, K, product=None, mof_both=True):
    """Faults of one scenario, one trial per vector.  mode: SBF, DBF, SBF+SBF, SBF+DBF, MOF1 (one element: kind-2 faults
    on BOTH products, apply_mof_to_signatures), MOF2.  product: None = random choice of mul1 / mul2 per fault (SBF / DBF,
    apply_sbf_or_dbf_to_signatures :157-168), or 1 / 2 to force it; mof_both = False: a MOF replaces the low bits of ONE
    product (outer_check.cpp's RAND op on one word, :160-166)."""
    parts = mode.split("+") if mode not in ("MOF1", "MOF2") else (["MOF"] if mode == "MOF1" else ["MOF", "MOF"])
    i1 = rng.integers(0, vec_len, size=trials)
    idx = [i1]
    if len(parts) == 2:
        i2 = rng.integers(0, vec_len - 1, size=trials)
        idx.append(np.where(i2 >= i1, i2 + 1, i2))
    out = []
    base = np.arange(trials, dtype=np.uint64) * np.uint64(vec_len)
    for part, ii in zip(parts, idx):
        if part == "MOF":
            for prod in ((1, 2) if mof_both else (product or 1,)):
                f = np.zeros(trials, dtype=FAULT_DT)
                f["index"] = base + ii.astype(np.uint64); f["product"] = prod; f["kind"] = 2
                f["payload"] = rng.integers(0, 2 ** 63, size=trials, dtype=np.uint64)
                out.append(f)
        else:
            f = np.zeros(trials, dtype=FAULT_DT)
            f["index"] = base + ii.astype(np.uint64); f["kind"] = 1
            f["product"] = rng.integers(1, 3, size=trials) if product is None else product
            f["payload"] = _bits(rng, trials, K, 1 if part == "SBF" else 2)
            out.append(f)
    return np.ascontiguousarray(np.concatenate(out))


def within_3_sigma(k1, n1, k2, n2):
    p = (k1 + k2) / (n1 + n2)
    sd = (p * (1 - p) * (1.0 / n1 + 1.0 / n2)) ** 0.5
    return abs(k1 / n1 - k2 / n2) <= 3.0 * sd + 1.0 / min(n1, n2)


def missed_signature(clean, faulty, trials):
    """barrett_final's verdict: a trial is detected when any fold group's accumulated signature differs (:259-264)"""
    return int(trials - (clean.reshape(trials, -1) != faulty.reshape(trials, -1)).any(axis=1).sum())


def missed_verdict(v, trials, bit):
    return int(trials - ((v.reshape(trials, -1) >> bit) & 1).any(axis=1).sum())
