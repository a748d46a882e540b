#!/usr/bin/env python
"""Config C5: key-switch ModUp (INTT -> HPS base conversion -> NTT), N=2^17, L=48 (+16 special
primes), dnum=3, limb-sharded over the GPUs of one box with ONE NCCL all-gather per digit.

  python bench_modup.py                                   # 1 GPU, no collective
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
         --master-port 29511 bench_modup.py               # limb-sharded, NCCL over NVLink

Every rank first checks its owned output limbs bit-for-bit against the unsharded (world = 1)
pipeline run on its own GPU, then the sharded ModUp of all dnum digits is timed with CUDA events
(max over ranks).  Prints one JSON line on rank 0.  Secondary to bench.py (the headline metric).
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log-n", type=int, default=17)
    ap.add_argument("--limbs", type=int, default=48)
    ap.add_argument("--special", type=int, default=16)
    ap.add_argument("--dnum", type=int, default=3)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--no-graph", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    import torch
    import torch.distributed as dist
    import fhe_reliability_gpu_b200 as pkg
    from fhe_reliability_gpu_b200 import modup
    if not torch.cuda.is_available():
        raise SystemExit("bench_modup.py needs CUDA devices (no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    log_n, N, n_q, n_p, dnum = args.log_n, 1 << args.log_n, args.limbs, args.special, args.dnum
    moduli = pkg.primes_create(N, [60] * (n_q + n_p))
    rng = np.random.default_rng(20250606)                      # identical on every rank
    a = np.stack([rng.integers(0, q, size=N, dtype=np.uint64) for q in moduli[:n_q]])

    plan = modup.ShardPlan(n_q, n_p, dnum, world, rank)
    be = modup.CudaBackend(pkg, log_n, moduli, rank, world, local)
    mine = {g: torch.from_numpy(a[g].view(np.int64)).cuda() for g in range(n_q) if g % world == rank}

    # ---- parity gate: sharded result == unsharded result on the limbs this rank owns
    ok = 1
    if world > 1:
        plan1 = modup.ShardPlan(n_q, n_p, dnum, 1, 0)
        be1 = modup.CudaBackend(pkg, log_n, moduli, 0, 1, local)
        all_limbs = {g: torch.from_numpy(a[g].view(np.int64)).cuda() for g in range(n_q)}
    for d in range(dnum):
        res = modup.modup_digit(plan, be, dist if world > 1 else None, d, mine)
        if world > 1:
            ref = modup.modup_digit(plan1, be1, None, d, all_limbs)
            for g, t in res.items():
                if not torch.equal(t, ref[g]):
                    ok = 0
    flag = torch.tensor([ok], device="cuda")
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if int(flag.item()) != 1:
        raise SystemExit("sharded ModUp differs from the unsharded pipeline")

    def step():
        for d in range(dnum):
            modup.modup_digit(plan, be, dist if world > 1 else None, d, mine)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    # the step is ~30 short launches: on one GPU replay it as one CUDA graph when capture succeeds (nothing in the
    # C ABI calls synchronises or allocates on this path after warm-up)
    graphed = False
    if not args.no_graph and world == 1:      # with NCCL inside, capture across ranks stalled on this stack: eager there
        try:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                step()
            torch.cuda.current_stream().wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                step()
            eager_step, step = step, graph.replay
            graphed = True
            step()
            barrier()
        except Exception as ex:                                   # capture not possible: stay eager, say so
            sys.stderr.write("CUDA graph capture failed, timing eager launches: %r\n" % (ex,))
            torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / args.steps
    alpha = n_q // dnum
    k_out = n_q + n_p - alpha
    if rank == 0:
        print(json.dumps({
            "metric": "ModUp (INTT + base conversion + NTT) polynomial raises/s", "value": 1e3 / ms, "unit": "polys/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms,
            "higher_is_better": True, "scaling": "strong", "cuda_graph": graphed, "dtype": "u64", "data": "synthetic",
            "config": {"workload": "C5: ModUp of one polynomial, N=2^%d, L=%d, K=%d special primes, dnum=%d, limb-sharded (limb g on rank g %% world)"
                       % (log_n, n_q, n_p, dnum), "collective": "one all_gather_into_tensor of %d scaled limbs per digit" % alpha if world > 1 else "none"},
            "work_per_step": {"limb_transforms": dnum * (alpha + k_out), "bconv_mulmods": dnum * alpha * (1 + k_out) * N,
                              "allgather_bytes": dnum * alpha * N * 8 if world > 1 else 0},
            "parity": "sharded == unsharded on every owned limb (bit-exact)" if world > 1 else "world 1 (tests/test_bconv_gpu.py checks it against the oracle)"}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
