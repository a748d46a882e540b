import sys, os
sys.path.insert(0, os.getcwd())
import torch, numpy as np
import fhe_reliability_gpu_b200 as pkg
from bench_ops import timed
log_n, alpha, K = 17, 16, 48
N = 1 << log_n
moduli = pkg.primes_create(N, [60] * (alpha + K))
ctx = pkg.Context(log_n, moduli)
conv = pkg.BaseConv(ctx, list(range(alpha)), list(range(alpha, alpha + K)))
g = torch.Generator(device="cuda").manual_seed(1)
y = torch.randint(0, 1 << 59, (alpha, N), dtype=torch.int64, device="cuda", generator=g)
o = torch.zeros((K, N), dtype=torch.int64, device="cuda")
sp = torch.cuda.current_stream().cuda_stream
for kc in (4, 8, 16, 24, 32, 48):
    ms = timed(torch, lambda: conv.contract(y.data_ptr(), o.data_ptr(), 0, kc, stream=sp), 50)
    print("kcount", kc, "ms %.4f" % ms)
