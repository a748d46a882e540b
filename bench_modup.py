#!/usr/bin/env python
"""Config C5: key-switch ModUp (INTT -> HPS base conversion -> NTT), N=2^17, L=48 (+16 special primes), dnum=3,
limb-sharded over the GPUs of one box through the C ABI (b200fhe_shard_* / b200fhe_modup_sharded: one NCCL all-gather per
digit on the plan's own stream, overlapped with the previous digit's contraction + NTT).

  python bench_modup.py                                   # 1 GPU, no collective
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
         --master-port 29511 bench_modup.py               # limb-sharded, NCCL over NVLink

Every rank first checks its owned output limbs bit-for-bit against the unsharded pipeline run on its own GPU, then the
sharded ModUp is timed with CUDA events (max over ranks).  Prints one JSON line on rank 0 (it contains "bit-exact" when the
parity gate passed).  Secondary to bench.py, whose line carries the same record as `modup_sharded` when WORLD_SIZE > 1.
torch.distributed is used for the rendezvous (handing rank 0's NCCL id to the other ranks) and the timing reductions only.
"""
import argparse
import json
import os
import sys

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
    import bench_sub
    if not torch.cuda.is_available():
        raise SystemExit("bench_modup.py needs CUDA devices (no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    peak = 6542.4
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peak = float(json.load(f)["hbm_gbs"])
    except Exception:
        pass
    rec = bench_sub.modup_sharded_record(pkg, torch, dist, rank, world, local, peak, args.log_n, args.limbs, args.special, args.dnum,
                                         args.steps, args.warmup, graph=not args.no_graph)
    if not rec["parity_vs_unsharded"]:
        raise SystemExit("sharded ModUp differs from the unsharded pipeline")
    if rank == 0:
        rec["parity"] = "sharded == unsharded on every owned limb (bit-exact)"
        rec["metric"] = "ModUp (INTT + base conversion + NTT) polynomial raises/s"
        print(json.dumps(rec))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
