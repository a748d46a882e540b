"""Profiling driver: forward / inverse transform of bench.py's C2 batch (8 polys x 36 limbs, N = 2^16), unprotected and protected,
`iters` times.  Every iteration launches, in this order: k_fwd_col, k_fwd_row | k_inv_row, k_inv_col | k_fwd_col<PROT>, k_fwd_row<PROT>,
k_abft_finalize | k_inv_row<PROT>, k_inv_col<PROT>, k_abft_finalize  (10 launches), so `ncu -k regex:'k_fwd|k_inv|k_abft' --launch-skip 10
--launch-count 10` captures one warm iteration.  Without ncu it prints CUDA-event times per call."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
import fhe_reliability_gpu_b200 as pkg

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 2
log_n, L, P = 16, 36, 8
N = 1 << log_n
moduli = pkg.primes_create(N, [60] * L)
ctx = pkg.Context(log_n, moduli)
rng = np.random.default_rng(1)
a = np.stack([np.stack([rng.integers(0, q, N, dtype=np.uint64) for q in moduli]) for _ in range(P)])
d = torch.from_numpy(a.view(np.int64)).cuda()
verdict = torch.zeros(P * L, dtype=torch.int32, device="cuda")
lib, h = pkg.lib(), ctx._h
s = torch.cuda.current_stream().cuda_stream
calls = [("forward", lambda: lib.b200fhe_nwt_forward(h, d.data_ptr(), P, L, 0, s)),
         ("inverse", lambda: lib.b200fhe_nwt_backward(h, d.data_ptr(), P, L, 0, s)),
         ("forward protected", lambda: lib.b200fhe_nwt_forward_protected(h, d.data_ptr(), P, L, 0, verdict.data_ptr(), None, None, 0, s)),
         ("inverse protected", lambda: lib.b200fhe_nwt_backward_protected(h, d.data_ptr(), P, L, 0, verdict.data_ptr(), None, s))]
for it in range(iters):
    for name, f in calls:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        assert f() == 0, lib.b200fhe_last_error()
        e1.record()
        torch.cuda.synchronize()
        if it == iters - 1:
            print("%-18s %.4f ms" % (name, e0.elapsed_time(e1)))
assert int(verdict.sum().item()) == 0
assert np.array_equal(d.cpu().numpy().view(np.uint64), a)
print("round trips exact, no fault flagged")
