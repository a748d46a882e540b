#!/usr/bin/env python
"""Secondary measurements (one JSON line each): config C3 pmult-accumulate and the two base-conversion
kernels of config C5, device-resident, CUDA events, against the rooflines DESIGN.md names.
bench.py stays the headline (NTT limb-transforms/s)."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def timed(torch, fn, steps, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    args = ap.parse_args()
    import torch
    import fhe_reliability_gpu_b200 as pkg
    if not torch.cuda.is_available():
        raise SystemExit("bench_ops.py needs a CUDA device (no CPU fallback)")
    peak = 6650.0
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = float(json.load(open(p))["hbm_gbs"])
    sp = torch.cuda.current_stream().cuda_stream

    # ---- C3: pmult, N=2^16, L=36, 64 terms, ciphertexts of 2 polynomials (3.41 GiB streamed per call)
    log_n, L, T, P = 16, 36, 64, 2
    N = 1 << log_n
    moduli = pkg.primes_create(N, [60] * L)
    ctx = pkg.Context(log_n, moduli)
    g = torch.Generator(device="cuda").manual_seed(20250606)
    ct = torch.randint(0, 1 << 59, (T, P, L, N), dtype=torch.int64, device="cuda", generator=g)
    pt = torch.randint(0, 1 << 59, (T, L, N), dtype=torch.int64, device="cuda", generator=g)
    out = torch.zeros((P, L, N), dtype=torch.int64, device="cuda")
    cps = [ct[k].data_ptr() for k in range(T)]
    pps = [pt[k].data_ptr() for k in range(T)]
    ms = timed(torch, lambda: ctx.pmult_accum(cps, pps, out.data_ptr(), P, L=L, stream=sp), args.steps)
    # spot parity on one limb against exact integer arithmetic
    l = 7
    want = np.zeros((P, 64), dtype=object)
    cth = ct[:, :, l, :64].cpu().numpy().astype(object)
    pth = pt[:, l, :64].cpu().numpy().astype(object)
    for k in range(T):
        want += cth[k] * pth[k][None, :]
    want %= moduli[l]
    got = out[:, l, :64].cpu().numpy().astype(object)
    assert (got == want).all(), "pmult parity failed"
    algo = (3 * T + 2) * L * N * 8
    print(json.dumps({"metric": "pmult ciphertext-term products/s", "value": T / (ms * 1e-3), "unit": "ct*pt terms/s", "ms_per_step": ms,
                      "config": {"workload": "C3: sum of 64 ct x pt, N=2^16, L=36, 2 polys per ct"}, "dtype": "u64",
                      "roofline": {"bound": "hbm", "achieved": algo / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                                   "frac": algo / (ms * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_launch": algo, "traffic": None},
                      "mulmods_per_s": 2 * T * L * N / (ms * 1e-3)}))
    del ct, pt, out

    # ---- evaluator steps around the key-switch (f3): tensor product and Galois permutation, N=2^16, L=36
    from fhe_reliability_gpu_b200.evaluator import Evaluator
    a2 = torch.randint(0, 1 << 59, (2, L, N), dtype=torch.int64, device="cuda", generator=g)
    b2 = torch.randint(0, 1 << 59, (2, L, N), dtype=torch.int64, device="cuda", generator=g)
    o3 = torch.zeros((3, L, N), dtype=torch.int64, device="cuda")
    o2 = torch.zeros((2, L, N), dtype=torch.int64, device="cuda")
    lib = pkg.lib()
    ms_t = timed(torch, lambda: pkg.binding.check(lib.b200fhe_ct_tensor(ctx._h, a2.data_ptr(), b2.data_ptr(), o3.data_ptr(), L, 0, sp)), args.steps)
    ms_g = timed(torch, lambda: pkg.binding.check(lib.b200fhe_galois_ntt(ctx._h, a2.data_ptr(), o2.data_ptr(), 5, 2 * L, sp)), args.steps)
    ms_a = timed(torch, lambda: pkg.binding.check(lib.b200fhe_poly_add(ctx._h, o2.data_ptr(), a2.data_ptr(), 2, L, 0, sp)), args.steps)
    w = L * N * 8
    print(json.dumps({"metric": "evaluator element-wise steps", "unit": "GB/s algorithmic", "dtype": "u64",
                      "config": {"workload": "f3: ct x ct tensor (4 in, 3 out), Galois permutation (2 in, 2 out), add_inplace (2+2 in, 2 out), N=2^16, L=36"},
                      "ct_tensor": {"ms": ms_t, "GBs": 7 * w / (ms_t * 1e-3) / 1e9, "frac_of_hbm_peak": 7 * w / (ms_t * 1e-3) / 1e9 / peak},
                      "galois_ntt": {"ms": ms_g, "GBs": 4 * w / (ms_g * 1e-3) / 1e9, "frac_of_hbm_peak": 4 * w / (ms_g * 1e-3) / 1e9 / peak},
                      "poly_add": {"ms": ms_a, "GBs": 6 * w / (ms_a * 1e-3) / 1e9, "frac_of_hbm_peak": 6 * w / (ms_a * 1e-3) / 1e9 / peak},
                      "roofline": {"bound": "hbm", "peak": peak, "unit": "GB/s",
                                   "note": "working sets of 75-130 MB sit partly in the 126 MB L2, so fractions above the HBM copy peak are L2 hits"}}))
    del a2, b2, o3, o2

    # ---- C5 kernels: one digit of ModUp, N=2^17, alpha=16 source limbs -> 48 target limbs
    log_n, alpha, K = 17, 16, 48
    N = 1 << log_n
    moduli = pkg.primes_create(N, [60] * (alpha + K))
    ctx2 = pkg.Context(log_n, moduli)
    conv = pkg.BaseConv(ctx2, list(range(alpha)), list(range(alpha, alpha + K)))
    x = torch.randint(0, 1 << 59, (alpha, N), dtype=torch.int64, device="cuda", generator=g)
    y = torch.zeros_like(x)
    o = torch.zeros((K, N), dtype=torch.int64, device="cuda")
    ms_s = timed(torch, lambda: conv.scale(x.data_ptr(), y.data_ptr(), 0, alpha, stream=sp), args.steps)
    ms_c = timed(torch, lambda: conv.contract(y.data_ptr(), o.data_ptr(), 0, K, stream=sp), args.steps)
    ver = torch.zeros(alpha + K, dtype=torch.int32, device="cuda")
    ms_u = timed(torch, lambda: conv.modup(x.data_ptr(), o.data_ptr(), stream=sp), args.steps)
    ms_p = timed(torch, lambda: conv.modup_protected(x.data_ptr(), o.data_ptr(), ver.data_ptr(), None, stream=sp), args.steps)
    assert int(ver.abs().sum().item()) == 0, "protected conversion flagged a fault-free run"
    algo = (alpha + K) * N * 8
    mm = alpha * (1 + K) * N
    print(json.dumps({"metric": "ModUp base-conversion digits/s", "value": 1e3 / (ms_s + ms_c), "unit": "digits/s",
                      "ms_scale": ms_s, "ms_contract": ms_c, "dtype": "u64",
                      "checksum_carrying": {"ms_unprotected": ms_u, "ms_protected": ms_p, "overhead_pct": 100.0 * (ms_p / ms_u - 1.0),
                                            "checks": "scale, y in memory, contraction: per limb, sums accumulated inside the conversion kernels"},
                      "config": {"workload": "C5 kernels: one digit, N=2^17, alpha=16 -> 48 target limbs (HPS)"},
                      "roofline": {"bound": "tensor pipe (byte-split mma.sync u8, 8 cycles per m16n8k32 per sub-partition) + integer epilogue (one 128-bit Barrett per output)",
                                   "imma_per_launch": (N // 16) * (K * 2) * ((alpha + 3) // 4), "tensor_pipe_frac": (N // 16) * (K * 2) * ((alpha + 3) // 4) * 8 / (592 * 1.965e9) / (ms_c * 1e-3),
                                   "mulmods_per_s": mm / ((ms_s + ms_c) * 1e-3),
                                   "hbm_achieved_GBs": algo / ((ms_s + ms_c) * 1e-3) / 1e9, "hbm_peak_GBs": peak,
                                   "algorithmic_bytes_per_digit": algo}}))


if __name__ == "__main__":
    main()
