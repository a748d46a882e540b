#!/usr/bin/env python
"""Row f1: the complete hybrid key-switch (INTT -> ModUp -> NTT -> sum_d up_d (.) evk_d -> INTT -> ModDown -> NTT) of one
polynomial at config C5's shape (N=2^17, 48 ciphertext limbs + 16 special primes, dnum=3), device-resident, one GPU.

  python bench_keyswitch.py [--log-n 17 --limbs 48 --special 16 --dnum 3 --steps 50]

Parity gate before timing: a real RLWE switching key is generated on the host (oracle arithmetic, test-side code) and
out0 + out1 s - c s' must come out as one small integer polynomial; then the switch is timed with CUDA events, eagerly
and as one CUDA-graph replay, plus its three pieces.  One JSON line.  Secondary to bench.py (the headline metric)."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


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
    ap.add_argument("--log-n", type=int, default=17)
    ap.add_argument("--limbs", type=int, default=48)
    ap.add_argument("--special", type=int, default=16)
    ap.add_argument("--dnum", type=int, default=3)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--no-check", action="store_true")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    import fhe_reliability_gpu_b200 as pkg
    if not torch.cuda.is_available():
        raise SystemExit("bench_keyswitch.py needs a CUDA device (no CPU fallback)")
    # under torchrun: one independent key-switch stream per GPU (ciphertext-parallel, no collective: weak scaling);
    # rank 0 alone runs the parity gate and the CPU baseline, the timed region is bracketed by barriers, max over ranks
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        if rank:
            args.no_check = True
    log_n, N, n_q, n_p, dnum = args.log_n, 1 << args.log_n, args.limbs, args.special, args.dnum
    total, alpha = n_q + n_p, n_q // dnum
    moduli = pkg.primes_create(N, [60] * total)
    ctx = pkg.Context(log_n, moduli, device=local)
    ks = pkg.KeySwitch(ctx, n_q, n_p, dnum)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).cuda()
    parity = "skipped (--no-check)"
    if not args.no_check:
        from oracle import oracle as O          # checker only
        from ks_common import check_switch_noise, make_switch_key
        tabs = [O.Tables(log_n, q) for q in moduli]
        rng = np.random.default_rng(20250606)
        s, s_new = rng.integers(-1, 2, N), rng.integers(-1, 2, N)
        evk = make_switch_key(O, rng, s, s_new, moduli, n_q, n_p, dnum, tabs)
        c = O.uniform_limbs(moduli[:n_q], N)
    else:
        g = np.random.default_rng(1)
        evk = g.integers(0, 1 << 59, (dnum, 2, total, N), dtype=np.uint64)
        c = g.integers(0, 1 << 59, (n_q, N), dtype=np.uint64)
    d_c, d_eq, d_ep = dev(c), dev(evk[:, :, :n_q]), dev(evk[:, :, n_q:])
    d_out = torch.zeros((2, n_q, N), dtype=torch.int64, device="cuda")
    sp = torch.cuda.current_stream().cuda_stream
    launches0 = pkg.lib().b200fhe_launch_count()
    ks.apply(d_c.data_ptr(), d_eq.data_ptr(), d_ep.data_ptr(), d_out.data_ptr(), stream=sp)
    torch.cuda.synchronize()
    launches = int(pkg.lib().b200fhe_launch_count() - launches0)
    cpu_baseline = None
    if not args.no_check:
        import time
        t0 = time.perf_counter()
        want = O.keyswitch_hybrid(c, evk, moduli, n_q, n_p, dnum, tabs)          # the CPU restatement, timed beside the GPU path
        dt = time.perf_counter() - t0
        if not np.array_equal(want, d_out.cpu().numpy().view(np.uint64)):
            raise SystemExit("key-switch differs from the oracle composition")
        cpu_baseline = {"value": 1.0 / dt, "unit": "key-switches/s", "cores": os.cpu_count(), "kind": "port",
                        "sample": "one key-switch of the same shape (oracle: OpenMP NTT batches + C base conversion / multiply-accumulate)"}
        noise = check_switch_noise(O, d_out.cpu().numpy().view(np.uint64), c, s, s_new, moduli[:n_q], tabs[:n_q])
        parity = "out0 + out1 s - c s' is one integer polynomial with max |coeff| = %d (RLWE key, sigma 3.2)" % int(np.abs(noise).max())

    step = lambda: ks.apply(d_c.data_ptr(), d_eq.data_ptr(), d_ep.data_ptr(), d_out.data_ptr(), stream=torch.cuda.current_stream().cuda_stream)
    ms_eager = timed(torch, step, args.steps)
    ms_graph = None
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            step()
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            step()
        ref = d_out.clone()
        d_out.zero_()
        graph.replay()
        torch.cuda.synchronize()
        if not torch.equal(ref, d_out):
            raise RuntimeError("graph replay differs from the eager result")
        ms_graph = timed(torch, graph.replay, args.steps)
    except Exception as ex:
        sys.stderr.write("CUDA graph capture failed, eager timing only: %r\n" % (ex,))
        torch.cuda.synchronize()

    verdict = torch.zeros((dnum + 2, total), dtype=torch.int32, device="cuda")
    ms_prot = timed(torch, lambda: ks.apply_protected(d_c.data_ptr(), d_eq.data_ptr(), d_ep.data_ptr(), d_out.data_ptr(), verdict.data_ptr(),
                                                      stream=torch.cuda.current_stream().cuda_stream), args.steps)
    if int(verdict.abs().sum().item()) != 0:
        raise SystemExit("protected key-switch flagged a fault-free run")
    vntt = torch.zeros(ks.ntt_count(), dtype=torch.int32, device="cuda")
    ms_full = timed(torch, lambda: ks.apply_protected_full(d_c.data_ptr(), d_eq.data_ptr(), d_ep.data_ptr(), d_out.data_ptr(), verdict.data_ptr(),
                                                           vntt.data_ptr(), stream=torch.cuda.current_stream().cuda_stream), args.steps)
    if int(verdict.abs().sum().item()) != 0 or int(vntt.abs().sum().item()) != 0:
        raise SystemExit("fully protected key-switch flagged a fault-free run")
    # pieces (eager)
    d_up = torch.zeros((dnum, total, N), dtype=torch.int64, device="cuda")
    accq = torch.zeros((2, n_q, N), dtype=torch.int64, device="cuda")
    accp = torch.zeros((2, n_p, N), dtype=torch.int64, device="cuda")
    ms_up = timed(torch, lambda: ks.modup(d_c.data_ptr(), d_up.data_ptr(), stream=sp), args.steps)
    ctq = [d_eq[d].data_ptr() for d in range(dnum)]; ptq = [d_up[d].data_ptr() for d in range(dnum)]
    ctp = [d_ep[d].data_ptr() for d in range(dnum)]; ptp = [d_up[d, n_q:].data_ptr() for d in range(dnum)]

    def inner():
        ctx.pmult_accum(ctq, ptq, accq.data_ptr(), 2, L=n_q, start_mod=0, stream=sp)
        ctx.pmult_accum(ctp, ptp, accp.data_ptr(), 2, L=n_p, start_mod=n_q, stream=sp)
    ms_ip = timed(torch, inner, args.steps)
    ms_down = timed(torch, lambda: ks.moddown(accq.data_ptr(), accp.data_ptr(), d_out.data_ptr(), 2, stream=sp), args.steps)

    ms = ms_graph if ms_graph is not None else ms_eager
    if world > 1:                                              # all ranks together: barrier, K steps each, barrier; max over ranks
        fn = graph.replay if ms_graph is not None else step
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        tmax = torch.tensor([e0.elapsed_time(e1) / args.steps], device="cuda", dtype=torch.float64)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        ms = float(tmax.item())
        if rank:
            dist.destroy_process_group()
            return
    transforms = n_q + dnum * (total - alpha) + 2 * n_p + 2 * n_q
    mulmods = dnum * alpha * (1 + total - alpha) * N + 2 * n_p * (1 + n_q) * N
    peak = 6650.0
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = float(json.load(open(p))["hbm_gbs"])
    # compulsory HBM traffic of the pipeline as built: every kernel reads its inputs and writes its outputs once
    words = (2 * n_q                                   # copy c -> x
             + 2 * n_q                                 # digit pass-through copies
             + 2 * n_q                                 # INTT (two passes would be 4; counted once = algorithmic)
             + dnum * (2 * alpha)                      # scale
             + dnum * (alpha + total - alpha)          # contract
             + dnum * 2 * (total - alpha)              # NTT of the raised limbs
             + dnum * 3 * total + 2 * total            # inner product: evk (2) + up (1) per digit, 2 outputs
             + 2 * 2 * n_p + 2 * 2 * n_p + 2 * (n_p + n_q) + 2 * 2 * n_q + 3 * 2 * n_q) * N
    print(json.dumps({
        "metric": "hybrid key-switches/s (one polynomial -> two)", "value": world * 1e3 / ms, "unit": "key-switches/s", "n_gpus": world, "scaling": "weak",
        "steps": args.steps, "ms_per_step": ms, "ms_eager": ms_eager, "ms_graph": ms_graph, "higher_is_better": True,
        "dtype": "u64", "data": "synthetic", "gpu_launches_per_step": launches,
        "config": {"workload": "f1: hybrid key-switch, N=2^%d, %d ciphertext limbs + %d special primes, dnum=%d, 60-bit" % (log_n, n_q, n_p, dnum)},
        "checksum_carrying_conversions": {"ms_eager": ms_prot, "overhead_pct_vs_eager": 100.0 * (ms_prot / ms_eager - 1.0),
                                          "checks": (dnum + 2) * total},
        "fully_protected": {"ms_eager": ms_full, "overhead_pct_vs_eager": 100.0 * (ms_full / ms_eager - 1.0),
                            "checks": (dnum + 2) * total + int(vntt.numel()),
                            "note": "conversion checks + one fused-checksum verdict per limb-transform (b200fhe_keyswitch_apply_protected_full)"},
        "pieces_ms": {"modup": ms_up, "evk_inner_product": ms_ip, "moddown": ms_down},
        "work_per_step": {"limb_transforms": transforms, "bconv_mulmods": mulmods, "algorithmic_bytes": words * 8},
        "roofline": {"bound": "integer multiplier pipe (NTT butterflies + dense base-conversion contraction); HBM shown for scale",
                     "hbm_achieved_GBs": words * 8 / (ms * 1e-3) / 1e9, "hbm_peak_GBs": peak,
                     "limb_transforms_per_s": transforms / (ms * 1e-3)},
        "parity": parity + ("; bit-exact against the oracle composition" if cpu_baseline else ""), "cpu_baseline": cpu_baseline}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
