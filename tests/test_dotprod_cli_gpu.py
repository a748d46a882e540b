"""python -m fhe_reliability_gpu_b200.dotprod: the dotprod_test experiment (reliability_test/dotprod_test.cu:67-228) with the
reference's CLI and report lines; fault-free the encrypted dot product is exact, injected flips are announced on
stderr in the reference's format and corrupt (essentially) every symbol."""
import io
import re

import pytest

pytestmark = pytest.mark.gpu


def _run(bits, symbols, seed, log_n=11):
    from fhe_reliability_gpu_b200 import dotprod
    out, err = io.StringIO(), io.StringIO()
    r = dotprod.run(bits, symbols, seed=seed, log_n=log_n, out=out, err=err)
    return r, out.getvalue(), err.getvalue()


def test_fault_free_run_matches_cpu_result():
    r, out, err = _run(1, 0, seed=1)
    assert r["match"] and r["symbol_errors"] == 0 and err == ""
    for line in ("Example: BGV HomMul test", "Plaintext matrix row size: 1024", "Compute x * y homomorphically...",
                 "Elementwise symbol errors: 0 / 2048", "Elementwise Hamming distance (bit errors): 0",
                 "Dot product bit errors (Hamming distance): 0", "Absolute difference   = 0", "✔ Dot product matches CPU result."):
        assert line in out, line
    assert re.search(r"Decrypted dot product = (\d+)\nExpected \(CPU\)         = \1\n", out)


@pytest.mark.parametrize("bits,symbols,seed", [(1, 1, 3), (2, 3, 4)])
def test_injected_flips_are_reported_and_propagate(bits, symbols, seed):
    r, out, err = _run(bits, symbols, seed)
    flips = re.findall(r"Injected bitflip @ idx=(\d+), bit=(\d+)", err)
    assert len(flips) == bits * symbols and all(int(b) < 64 and int(i) < 2 * 4 * 2048 for i, b in flips)
    m = re.search(r"Elementwise symbol errors: (\d+) / (\d+)", out)
    # a flip in a ciphertext word reaches every slot unless flips on the same word cancel (same bit drawn twice)
    distinct = {(i, b) for i, b in flips}
    if len(distinct) == len(flips):
        assert int(m.group(1)) > 0.99 * int(m.group(2)) and "✖ MISMATCH detected!" in out and not r["match"]


def test_full_size_parameters_fault_free():
    """the reference's parameters: N = 16384, six 50-bit primes (two special), 20-bit plain modulus"""
    r, out, _ = _run(1, 0, seed=9, log_n=14)
    assert r["match"] and r["symbol_errors"] == 0 and r["slot_count"] == 16384


def test_usage_error_exit_code():
    from fhe_reliability_gpu_b200 import dotprod
    assert dotprod.main(["7"]) == 1
