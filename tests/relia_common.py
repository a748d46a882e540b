"""Shared helpers of the relia_ntt_sim parity tests: the simulator's fault scenarios (relia_ntt_sim.cpp:401-479) drawn
with numpy instead of mt19937_64, and a parser for the table the reference binary prints."""
import os
import re
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SIM = os.path.join(ROOT, "oracle", "_ref", "relia_ntt_sim")
SBF, DBF, MLF = 1, 2, 3
SCENARIOS = [("SBF", [SBF]), ("DBF", [DBF]), ("MLF", [MLF]), ("SBF+SBF", [SBF, SBF]), ("SBF+DBF", [SBF, DBF]),
             ("MLF+SBF", [MLF, SBF]), ("MLF+DBF", [MLF, DBF]), ("MLF+MLF", [MLF, MLF])]
STAGES = ["Stage1(Batch1)", "Stage2(Twiddle)", "Stage3(Batch2)"]
FAULT_DT = np.dtype([("trial", "<u4"), ("kind", "<u4"), ("op", "<u8"), ("payload", "<u8")])


def draw_faults(rng, kinds, trials, lo, hi, bits):
    """One fault per entry of `kinds` and per trial, at distinct operations of [lo, hi) (the simulator's used_ops set),
    payloads as inject_one (:114-133): SBF one random bit below `bits`, DBF two distinct bits, MLF a random word."""
    nf = len(kinds)
    out = np.zeros(trials * nf, dtype=FAULT_DT)
    ops = np.empty((trials, nf), dtype=np.int64)
    ops[:, 0] = rng.integers(lo, hi, size=trials)
    for j in range(1, nf):
        o = rng.integers(lo, hi, size=trials)
        while True:
            clash = (o[:, None] == ops[:, :j]).any(axis=1)
            if not clash.any():
                break
            o[clash] = rng.integers(lo, hi, size=int(clash.sum()))
        ops[:, j] = o
    for j, kind in enumerate(kinds):
        sl = out[j::nf]
        sl["trial"] = np.arange(trials, dtype=np.uint32)
        sl["op"] = ops[:, j]
        if kind == MLF:
            sl["kind"] = 2
            sl["payload"] = rng.integers(0, 2 ** 64, size=trials, dtype=np.uint64)
        else:
            sl["kind"] = 1
            b1 = rng.integers(0, bits, size=trials)
            p = np.left_shift(np.uint64(1), b1.astype(np.uint64))
            if kind == DBF:
                b2 = rng.integers(0, bits - 1, size=trials)
                b2 = np.where(b2 >= b1, b2 + 1, b2)
                p = p | np.left_shift(np.uint64(1), b2.astype(np.uint64))
            sl["payload"] = p
    # the simulator applies a SORTED plan; entries of one trial are distinct operations, so order is immaterial
    return out


def run_simulator(N, q_bits, trials, fold_bits=4):
    """{(scenario, stage): dict(det=count, b1=.., intra=.., inter=.., b2=.., trials=..)} from the reference binary's table."""
    txt = subprocess.run([SIM, "--N", str(N), "--q-bits", str(q_bits), "--trials", str(trials), "--fold-bits", str(fold_bits)],
                         capture_output=True, text=True, timeout=1800, check=True).stdout
    q = int(re.search(r"Q=(\d+)", txt).group(1))
    table, scen = {}, None
    for line in txt.splitlines():
        m = re.match(r"Scenario: \[(.+)\]", line)
        if m:
            scen = m.group(1)
            continue
        m = re.match(r"\s*(Stage\d\(\w+\))\s+([\d.]+)\s+([\d.]+)\s+\|\s+(\d+)\s+(\d+)\s+(\d+)\s+(\d+)", line)
        if m and scen:
            table[(scen, m.group(1))] = dict(det=round(float(m.group(2)) * trials), b1=int(m.group(4)), intra=int(m.group(5)),
                                             inter=int(m.group(6)), b2=int(m.group(7)), trials=trials)
    return q, table


def within_3_sigma(k1, n1, k2, n2):
    """two-sided test that two binomial counts share one rate (pooled normal approximation, with a continuity allowance)"""
    p = (k1 + k2) / (n1 + n2)
    sd = (p * (1 - p) * (1.0 / n1 + 1.0 / n2)) ** 0.5
    return abs(k1 / n1 - k2 / n2) <= 3.0 * sd + 1.0 / min(n1, n2)


STD_SIM = os.path.join(ROOT, "oracle", "_ref", "non_four_step_ntt_sim")


def run_standard_simulator(N, q_bits, trials):
    """{scenario: dict(det=count, lazy=.., inter=.., trials=..)} from the table non_four_step_ntt_sim prints (:267-330)."""
    txt = subprocess.run([STD_SIM, "--N", str(N), "--q-bits", str(q_bits), "--trials", str(trials)],
                         capture_output=True, text=True, timeout=1800, check=True).stdout
    q = int(re.search(r"Q=(\d+)", txt).group(1))
    table = {}
    for line in txt.splitlines():
        m = re.match(r"\s*(\S+)\s+Standard NTT\s+([\d.]+)\s+([\d.]+)\s+\|\s+(\d+)\s+(\d+)", line)
        if m:
            table[m.group(1)] = dict(det=round(float(m.group(2)) * trials), lazy=int(m.group(4)), inter=int(m.group(5)), trials=trials)
    return q, table
