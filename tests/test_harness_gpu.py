"""Harness-level parity (SURVEY 8a row a8, 8b iii): the ntt_test / pmult_fault_test executables keep the
reference's CLI and text contract, and the fault-propagation laws the reference recorded hold.

Contract sources: reliability_test/ntt_test.cu:125-127,171-198 (lines printed),
test_scripts/gen_errorimpact.py:29-30 (the regexes the sweep scripts apply to stderr),
reliability_test/symbolerror.txt:1 / errorset.txt:1 (recorded outputs: a flipped word corrupts
EVERY output symbol of its limb), reliability_test/dotprod_test.cu:55-58,136-139.
"""
import os
import re
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NTT_TEST = os.path.join(ROOT, "harness", "bin", "ntt_test")
DOTPROD_TEST = os.path.join(ROOT, "harness", "bin", "pmult_fault_test")
REF_SHIM = os.path.join(ROOT, "oracle", "_ref", "ntt_test_shim")
REF_REAL_SHIM = os.path.join(ROOT, "oracle", "_ref", "ntt_real_test_shim")

# the two regexes of test_scripts/gen_errorimpact.py:29-30
RE_BIT = re.compile(r"Bit error: \d+/\d+ = ([0-9.]+)")
RE_SYM = re.compile(r"Symbol error: \d+/\d+ = ([0-9.]+)")
RE_FULL = re.compile(r"\[FAULT DETECTED\] Bit error: (\d+)/(\d+) = ([0-9.]+), Symbol error: (\d+)/(\d+) = ([0-9.]+)")


def run(exe, *args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([exe] + [str(a) for a in args], capture_output=True, text=True, timeout=300, env=e)


def check_report(p, log_dim, batch, flips, syms, exact_limbs=None):
    assert p.returncode == 0, p.stderr
    total = batch << log_dim
    assert "[2D] Flipping %d bits across %d symbols..." % (flips * syms, syms) in p.stdout
    m = RE_FULL.search(p.stderr)
    assert m, p.stderr
    assert RE_BIT.search(p.stderr) and RE_SYM.search(p.stderr)
    ham, bits, ber, bad, tot, ser = int(m[1]), int(m[2]), float(m[3]), int(m[4]), int(m[5]), float(m[6])
    assert bits == total * 64 and tot == total
    assert abs(ber - ham / bits) < 1e-6 and abs(ser - bad / tot) < 1e-6
    assert "ERROR! Total bitwise Hamming distance = %d (bit error rate = " % ham in p.stdout
    assert "       Affected symbols = %d/%d (symbol error rate = " % (bad, tot) in p.stdout
    # propagation law: each hit limb is corrupted in all N outputs, so S is a multiple of N ...
    N = 1 << log_dim
    assert bad % N == 0 and N <= bad <= min(syms, batch) * N
    if exact_limbs is not None:
        assert bad == exact_limbs * N
    # ... and about half of the ~50 significant bits of every corrupted word differ
    per_word = ham / bad
    assert 20.0 < per_word < 30.0, per_word          # flipimpact_ntt.csv records 25/64
    return bad


@pytest.mark.parametrize("shape", [(12, 32, 1, 1), (12, 32, 2, 7), (12, 1, 3, 5), (17, 1, 1, 1), (13, 4, 64, 2)])
def test_ntt_test_contract_and_propagation(shape):
    log_dim, batch, flips, syms = shape
    p = run(NTT_TEST, log_dim, batch, flips, syms, 12345)
    check_report(p, log_dim, batch, flips, syms, exact_limbs=1 if (syms == 1 or batch == 1) else None)
    if batch == 1:
        assert "[Debug] Symbol mismatch indices:" not in p.stdout      # more than 128 mismatches


def test_ntt_test_is_reproducible_with_a_seed_and_random_without():
    a = run(NTT_TEST, 12, 8, 2, 3, 99)
    b = run(NTT_TEST, 12, 8, 2, 3, 99)
    assert a.stdout == b.stdout and a.stderr == b.stderr
    c = run(NTT_TEST, 12, 8, 2, 3, env={"B200FHE_SEED": "99"})
    assert c.stderr == a.stderr


def test_ntt_test_argument_errors_exit_1():
    assert run(NTT_TEST, 12, 1).returncode == 1                         # ntt_test.cu:202-206 (3 arguments = the legacy CLI below)
    assert "Usage:" in run(NTT_TEST, 12, 1).stderr
    assert run(NTT_TEST, 12, 1, 0).returncode == 1
    assert run(NTT_TEST, 12, 1, 0, 1).returncode == 1                   # :213-216
    p = run(NTT_TEST, 12, 1, 65, 1)
    assert p.returncode == 1 and "exceeds bits per symbol (64)" in p.stderr       # :29-33
    p = run(NTT_TEST, 4, 1, 1, 17)
    assert p.returncode == 1 and "exceeds total symbols (16)" in p.stderr         # :34-38


@pytest.mark.parametrize("n", [2, 32, 2048])
def test_legacy_three_argument_cli(n):
    """reliability_test/run_bench_test.sh:10 `./ntt_test 12 32 $n` (n unique random bit flips anywhere in the buffer) and the
    lines exp_log.txt:3-8 records for it.  n = 2 hits two limbs: symbol errors 8192/131072, as exp_log.txt:3 shows."""
    p = run(NTT_TEST, 12, 32, n, env={"B200FHE_SEED": str(1000 + n)})
    assert p.returncode == 0, p.stderr
    lines = p.stdout.splitlines()
    assert lines[0] == "Doing NTT 2D RADIX8"
    assert lines[1] == "[2D] Flipping %d unique random bits in the data buffer..." % n
    assert lines[2] == "Doing NTT 2D RADIX8"
    m = RE_FULL.search(p.stderr)
    assert m, p.stderr
    ham, bits, bad, tot = int(m[1]), int(m[2]), int(m[4]), int(m[5])
    assert bits == 32 * 4096 * 64 and tot == 32 * 4096
    assert bad % 4096 == 0 and 4096 <= bad <= min(n, 32) * 4096           # every output symbol of each hit limb
    if n == 2:
        assert bad in (4096, 8192)
    if n == 2048:
        assert bad == 32 * 4096                                             # exp_log.txt: large n corrupts every limb
    assert 20.0 < ham / bad < 30.0
    assert "ERROR! Total bitwise Hamming distance = %d (bit error rate = " % ham in p.stdout


def test_real_test_mode_reports_all_correct():
    """ntt_real_test.cu (injection disabled, :108-117): two clean runs must agree."""
    p = run(NTT_TEST, 12, 32, 0, 0, 7, env={"B200FHE_REAL_TEST": "1"})
    assert p.returncode == 0 and "ALL CORRECT" in p.stdout and "FAULT" not in p.stderr


@pytest.mark.skipif(not os.path.exists(REF_SHIM), reason="reference harness binary not built (needs /root/reference at build time)")
@pytest.mark.parametrize("shape", [(12, 32, 1, 1), (12, 32, 1, 16), (17, 1, 2, 1)])
def test_reference_ntt_test_source_unmodified_on_the_shim(shape):
    """reliability_test/ntt_test.cu itself, compiled against shim/phantom.h: same report, same laws."""
    log_dim, batch, flips, syms = shape
    p = run(REF_SHIM, log_dim, batch, flips, syms)
    check_report(p, log_dim, batch, flips, syms, exact_limbs=1 if syms == 1 else None)


@pytest.mark.skipif(not os.path.exists(REF_REAL_SHIM), reason="reference harness binary not built")
def test_reference_ntt_real_test_source_on_the_shim():
    p = run(REF_REAL_SHIM, 12, 32, 1)
    assert p.returncode == 0, p.stderr
    assert "ALL CORRECT" in p.stdout


def test_pmult_fault_test_injection_report():
    p = run(DOTPROD_TEST, 2, 3, 12, 3, 8, 4242)
    assert p.returncode == 0, p.stderr
    inj = re.findall(r"Injected bitflip @ idx=(\d+), bit=(\d+)", p.stderr)          # dotprod_test.cu:57-58
    assert len(inj) == 6
    m = re.search(r"Elementwise symbol errors: (\d+) / (\d+)", p.stdout)             # :136-137
    h = re.search(r"Elementwise Hamming distance \(bit errors\): (\d+)", p.stdout)   # :138-139
    assert m and h
    # a pointwise product-accumulate confines each corrupted input word to ONE output word
    # (unless both flips of a word hit the same bit and cancel)
    words = {}
    for idx, bit in inj:
        words.setdefault(idx, set()).symmetric_difference_update({bit})
    live = sum(1 for v in words.values() if v)
    assert int(m[1]) == live and int(m[2]) == 2 * 3 * 4096
    assert ("MISMATCH detected" in p.stdout) == (live > 0)


def test_pmult_fault_test_clean_run_matches():
    p = run(DOTPROD_TEST, 0, 0, 12, 2, 4, 1)
    assert p.returncode == 0 and "Elementwise symbol errors: 0 / " in p.stdout and "matches" in p.stdout
