"""NTT / INTT device rate against N and limbs per launch (the key-switch runs the transforms at N = 2^17 in launches of
16-144 limbs, bench.py measures N = 2^16 in launches of 288)."""
import os, sys
sys.path.insert(0, os.getcwd())
import torch
import fhe_reliability_gpu_b200 as pkg
from bench_ops import timed

for log_n in (15, 16, 17):
    N = 1 << log_n
    moduli = pkg.primes_create(N, [60] * 64)
    ctx = pkg.Context(log_n, moduli)
    sp = torch.cuda.current_stream().cuda_stream
    for L in (16, 32, 48, 64):
        for npoly in (1, 3):
            x = torch.randint(0, 1 << 59, (npoly, L, N), dtype=torch.int64, device="cuda")
            f = timed(torch, lambda: ctx.nwt_forward(x.data_ptr(), npoly, L=L, start_mod=0, stream=sp), 30)
            b = timed(torch, lambda: ctx.nwt_backward(x.data_ptr(), npoly, L=L, start_mod=0, stream=sp), 30)
            n = npoly * L
            print("N=2^%d L=%2d polys=%d : fwd %.1f us (%.2f us/limb, %.2f M limb-transforms/s)  inv %.1f us (%.2f us/limb)"
                  % (log_n, L, npoly, f * 1e3, f * 1e3 / n, n / f / 1e3, b * 1e3, b * 1e3 / n))
