"""the BGV host roles behind the C ABI (b200fhe_bgv_*) and the Phantom class surface of shim/ on top.

  * through ctypes at N = 4096: encode / decode are inverse and follow SEAL's slot order (a row rotation by k moves slot
    i + k to slot i; the column rotation swaps the two rows), encrypt -> decrypt is exact, multiply + relinearize +
    mod_switch_to_next decrypts to the slot-wise product (with the correction factor), rotate-and-add gives the dot product;
  * the reference's OWN reliability_test/dotprod_test.cu, compiled UNMODIFIED against shim/ into oracle/_ref/dotprod_test_shim
    (harness/Makefile `ref`): fault-free it prints the reference's success line, after one injected bit flip every slot is
    wrong (16384 / 16384, the propagation law recorded in reliability_test/data/bits1-16_num1.txt:30)."""
import os
import re
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "dotprod_test_shim")


def _setup(pkg, log_n=12, seed=5):
    N = 1 << log_n
    moduli = pkg.primes_create(N, [50] * 6)
    t = pkg.primes_create(N, [20])[0]
    return pkg.binding.Bgv(log_n, moduli, 2, t, seed=seed), N, t


def test_encoder_slot_order_and_encrypt_decrypt_roundtrip(pkg):
    import torch
    bgv, N, t = _setup(pkg)
    rng = np.random.default_rng(1)
    v = rng.integers(0, t, N, dtype=np.uint64)
    plain = bgv.encode(v)
    assert (bgv.decode(plain) == v).all()
    # the constant vector encodes to the constant polynomial
    assert (bgv.encode(np.full(N, 7, dtype=np.uint64)) == np.array([7] + [0] * (N - 1), dtype=np.uint64)).all()
    ct = torch.zeros((2, bgv.n_q, N), dtype=torch.int64, device="cuda")
    bgv.encrypt(plain, ct.data_ptr())
    assert (bgv.decode(bgv.decrypt(ct.data_ptr(), 2, bgv.n_q)) == v).all()
    # rotations: rows left by k, and the row swap
    row = N // 2
    for k in (1, 5, row - 1):
        r = ct.clone()
        bgv.rotate(r.data_ptr(), bgv.n_q, k)
        got = bgv.decode(bgv.decrypt(r.data_ptr(), 2, bgv.n_q))
        want = np.concatenate([np.roll(v[:row], -k), np.roll(v[row:], -k)])
        assert (got == want).all(), k
    r = ct.clone()
    bgv.rotate(r.data_ptr(), bgv.n_q, 0, columns=True)
    assert (bgv.decode(bgv.decrypt(r.data_ptr(), 2, bgv.n_q)) == np.concatenate([v[row:], v[:row]])).all()
    bgv.close()


def test_multiply_relinearize_modswitch_rotate_sum(pkg):
    import torch
    bgv, N, t = _setup(pkg, seed=9)
    rng = np.random.default_rng(2)
    a, b = rng.integers(0, 100, N, dtype=np.uint64), rng.integers(0, 100, N, dtype=np.uint64)
    nq = bgv.n_q
    ca = torch.zeros((2, nq, N), dtype=torch.int64, device="cuda"); cb = torch.zeros_like(ca)
    bgv.encrypt(bgv.encode(a), ca.data_ptr()); bgv.encrypt(bgv.encode(b), cb.data_ptr())
    c3 = torch.zeros((3, nq, N), dtype=torch.int64, device="cuda")
    bgv.multiply(ca.data_ptr(), cb.data_ptr(), c3.data_ptr(), nq)
    assert (bgv.decode(bgv.decrypt(c3.data_ptr(), 3, nq)) == a * b % t).all()          # size-3 decryption
    bgv.relinearize(c3.data_ptr(), nq)
    c2 = c3[:2].contiguous()
    assert (bgv.decode(bgv.decrypt(c2.data_ptr(), 2, nq)) == a * b % t).all()
    lo = torch.zeros((2, nq - 1, N), dtype=torch.int64, device="cuda")
    cf = bgv.mod_switch_to_next(c2.data_ptr(), 2, nq, lo.data_ptr())
    assert cf != 1
    assert (bgv.decode(bgv.decrypt(lo.data_ptr(), 2, nq - 1, cf)) == a * b % t).all()
    assert not (bgv.decode(bgv.decrypt(lo.data_ptr(), 2, nq - 1, 1)) == a * b % t).all()   # without the correction factor it is m q_last^-1
    step, row = 1, N // 2
    while step < row:                                                                     # dotprod_test.cu:143-149
        r = lo.clone()
        bgv.rotate(r.data_ptr(), nq - 1, step)
        bgv.add(lo.data_ptr(), r.data_ptr(), 2, nq - 1)
        step <<= 1
    res = bgv.decode(bgv.decrypt(lo.data_ptr(), 2, nq - 1, cf))
    assert (int(res[0]) + int(res[row])) % t == int((a * b % t).sum()) % t
    # one flipped ciphertext bit corrupts every slot (NTT-form ciphertext, reliability_test/data/flipimpact_ntt.csv law)
    pkg.binding.flip_bit(ca.data_ptr(), 12345, 17)
    torch.cuda.synchronize()
    bad = bgv.decode(bgv.decrypt(ca.data_ptr(), 2, nq))
    assert (bad != a).sum() > 0.99 * N
    bgv.close()


def _run(args, seed):
    env = dict(os.environ, B200FHE_SEED=str(seed))
    return subprocess.run([REF_BIN] + args, capture_output=True, text=True, timeout=600, env=env)


@pytest.mark.skipif(not os.path.exists(REF_BIN), reason="oracle/_ref/dotprod_test_shim not built (reference tree absent at build time)")
def test_reference_dotprod_test_compiled_unmodified_runs_on_the_shim():
    p = _run(["0", "0"], 3)
    assert p.returncode == 0, p.stderr[-2000:]
    out = p.stdout
    assert "|   coeff_modulus size: 300 (50 + 50 + 50 + 50 + 50 + 50) bits" in out                       # data/bits1-16_num1.txt:8
    assert "1125899903107073 ,  1125899903500289 ,  1125899903795201 ,  1125899903827969 ,  1125899903991809 ,  1125899904679937 ,  " in out  # :10
    assert "Plaintext matrix row size: 8192" in out
    assert "Elementwise symbol errors: 0 / 16384" in out
    assert "✔ Dot product matches CPU result." in out
    m = re.search(r"Decrypted dot product = (\d+)\nExpected \(CPU\)         = (\d+)", out)
    assert m and m.group(1) == m.group(2)
    p = _run(["1", "1"], 4)
    assert p.returncode == 0, p.stderr[-2000:]
    assert re.search(r"Injected bitflip @ idx=\d+, bit=\d+", p.stderr)
    m = re.search(r"Elementwise symbol errors: (\d+) / 16384", p.stdout)                                    # :30 records 16384 / 16384
    assert m and int(m.group(1)) >= 16380, p.stdout[-600:]     # a corrupted slot still matches by chance with probability 1 / t each
    assert "MISMATCH detected!" in p.stdout or "✔ Dot product matches CPU result." in p.stdout
    assert _run([], 1).returncode == 1                                                                      # usage, exit 1 (:190-194)
