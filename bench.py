#!/usr/bin/env python
"""Headline benchmark: NTT limb-transforms/s at N=2^16, 60-bit RNS (BASELINE.json).

Workload (config C2, reliability_test/ntt_test.cu shape scaled to N=2^16, L=36):
a batch of NPOLY polynomials x 36 limbs, forward NTT of the whole batch then
inverse NTT of the whole batch, in place.  One step = 2*NPOLY*36 limb-transforms.
The batch (144 MiB at NPOLY=8) is larger than the 126 MB L2, so every transform
reads its input from HBM ("inputs larger than L2").

  value  : device-resident throughput, CUDA events, max over ranks
  e2e    : same metric through b200fhe_nwt_{forward,backward}_host with pinned
           HOST buffers (H2D + kernels + D2H inside the timed region)
  --impl reference : the CPU oracle (restated SEAL Harvey NTT, all host cores)
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LOG_N, L, BITS, NPOLY = 16, 36, 60, 8
N = 1 << LOG_N
SEED = 20250606
MIB = 1 << 20
ALGO_BYTES_PER_LIMB_TRANSFORM = 16 * N     # read once + write once


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def cpu_reference(steps, warmup, sample_polys=NPOLY):
    """Oracle (restated SEAL Harvey NTT) on ALL host cores: fwd+inv of sample_polys x 36 limbs per step, the same batch
    shape the GPU arm times.  Tables are stacked once (outside the timed loop); the OpenMP thread count is set explicitly
    (torchrun exports OMP_NUM_THREADS=1) and reported; the object is built with -march=native for this host."""
    os.environ.setdefault("ORACLE_NATIVE", "1")
    from oracle import oracle as O
    threads = O.set_threads(os.cpu_count() or 1)
    moduli = O.primes_create(N, [BITS] * L)
    bt = O.BatchTables([O.Tables(LOG_N, q) for q in moduli])
    a = np.stack([O.uniform_limbs(moduli, N, seed=SEED + i) for i in range(sample_polys)])
    ref = a.copy()
    def step():
        O.ntt_forward_batch(a, bt, inplace=True)
        O.ntt_inverse_batch(a, bt, inplace=True)
    for _ in range(warmup):
        step()
    if steps is None:                       # size the sample to about 10 s of CPU work
        t0 = time.perf_counter(); step(); one = time.perf_counter() - t0
        steps = int(min(2000, max(3, 10.0 / max(one, 1e-6))))
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = time.perf_counter() - t0
    assert np.array_equal(a, ref), "oracle round trip failed"
    per_step = 2 * sample_polys * L
    lib_name = os.path.basename(O.lib()._name)
    return per_step * steps / dt, dt / steps * 1e3, threads, ("%d polys x %d limbs fwd+inv per step (%d limb-transforms), %d steps, %d OpenMP threads, %s"
                                                                % (sample_polys, L, per_step, steps, threads, lib_name))


def newest_ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of each pass kernel, parsed from the newest committed
    profiles/*_ncu_full_summary.txt (ncu --set full of this same command; profiles/ncu_summary.py wrote it)."""
    import ast, glob, re
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_ncu_full_summary.txt")), key=lambda f: (os.path.getmtime(f), f))
    for f in reversed(files):
        rows = {}
        for line in open(f):
            m = re.match(r"(Kernel Name|dram__bytes_read\.sum|dram__bytes_write\.sum)\s+(\[.*\])\s*$", line)
            if m:
                rows[m.group(1)] = ast.literal_eval(m.group(2))
        if len(rows) == 3:
            out = {}
            for name, r, w in zip(rows["Kernel Name"], rows["dram__bytes_read.sum"], rows["dram__bytes_write.sum"]):
                for k in ("k_fwd_col", "k_fwd_row", "k_inv_row", "k_inv_col"):
                    if k in name and k not in out:                # the summary is in MB; the first launch of each name = the unprotected kernel
                        out[k] = (float(r) + float(w)) * 1e6
            if out:
                return out, os.path.relpath(f, ROOT)
    return {}, None


def measured_butterfly_cycles():
    """Best VALID register-only rate of the product's butterfly form (cycles per warp-butterfly per sub-partition) from the
    newest committed microbenchmark output under profiles/microbench/."""
    import glob, re
    best, src = None, None
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "microbench", "*butterfly*.txt")), key=lambda f: (os.path.getmtime(f), f))
    for f in reversed(files):
        for line in open(f):
            if line.startswith("V7") and "cyc/warp-butterfly/SMSP=" in line:      # V7 = the carry-chain form the product uses
                v = float(line.rsplit("=", 1)[1])
                if best is None or v < best:
                    best, src = v, os.path.relpath(f, ROOT)
        if best is not None:
            break
    return best, src


class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, dev):
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(dev), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.p = None

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        out = self.p.communicate()[0]
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except Exception:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=500)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--npoly", type=int, default=NPOLY)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sub", action="store_true", help="skip the pmult / ModUp / key-switch sub-records")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    config = {"workload": "C2: N=2^16, L=36 x 60-bit RNS limbs, batch of %d polys per GPU, forward NTT of the batch then inverse NTT of the batch, in place" % args.npoly,
              "log_n": LOG_N, "limbs": L, "prime_bits": BITS, "polys_per_gpu": args.npoly,
              "cache": "inputs larger than L2 (%d MiB per GPU vs 126 MB L2)" % (args.npoly * L * N * 8 // MIB),
              "sharding": "limb/ciphertext-sharded, no collective", "seed": SEED}

    if args.impl == "reference":
        if rank != 0:
            return
        steps = max(1, min(args.steps, 20)); warm = max(1, min(args.warmup, 3))
        v, ms, cores, sample = cpu_reference(steps, warm, args.npoly)
        print(json.dumps({"impl": "reference", "metric": "NTT limb-transforms/s", "value": v, "unit": "limb-transforms/s",
                          "n_gpus": args.gpus, "steps": steps, "warmup": warm, "ms_per_step": ms, "higher_is_better": True,
                          "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic", "config": config,
                          "cpu_baseline": {"value": v, "unit": "limb-transforms/s", "cores": cores, "kind": "port", "sample": sample},
                          "e2e": {"value": v, "unit": "limb-transforms/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    import torch
    import torch.distributed as dist
    import fhe_reliability_gpu_b200 as pkg
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = pkg.lib()
    moduli = pkg.primes_create(N, [BITS] * L)
    ctx = pkg.Context(LOG_N, moduli, device=local)
    npoly = args.npoly
    words = npoly * L * N
    # synthetic input: uniform residues, seeded per rank (disjoint data per GPU)
    rng = np.random.default_rng(SEED + rank)
    h = torch.empty(words, dtype=torch.int64).pin_memory()
    hv = h.numpy().view(np.uint64).reshape(npoly, L, N)
    for l, q in enumerate(moduli):
        hv[:, l, :] = rng.integers(0, q, size=(npoly, N), dtype=np.uint64)
    ref0 = hv[0].copy()
    d = h.to("cuda", non_blocking=False)
    stream = torch.cuda.current_stream()
    sp = stream.cuda_stream
    dp = d.data_ptr()

    def step():
        pkg.binding.check(lib.b200fhe_nwt_forward(ctx._h, dp, npoly, L, 0, sp))
        pkg.binding.check(lib.b200fhe_nwt_backward(ctx._h, dp, npoly, L, 0, sp))

    # correctness gate before timing: forward of poly 0 vs oracle, then round trip
    from oracle import oracle as O
    NOCHECK = bool(os.environ.get("B200_BENCH_NOCHECK"))   # timing ablations of deliberately wrong kernels
    if rank == 0 and not NOCHECK:          # every limb of two polynomials of the batch against the oracle, then the round trip
        ncheck = min(2, npoly)
        bt = O.BatchTables([O.Tables(LOG_N, q) for q in moduli])
        pkg.binding.check(lib.b200fhe_nwt_forward(ctx._h, dp, npoly, L, 0, sp))
        got = d[:ncheck * L * N].cpu().numpy().view(np.uint64).reshape(ncheck, L, N)
        assert np.array_equal(got, O.ntt_forward_batch(hv[:ncheck], bt)), "GPU forward != oracle"
        pkg.binding.check(lib.b200fhe_nwt_backward(ctx._h, dp, npoly, L, 0, sp))
        assert np.array_equal(d.cpu().numpy().view(np.uint64).reshape(npoly, L, N), hv), "round trip failed"

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    l0 = lib.b200fhe_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    barrier()
    launches = int(lib.b200fhe_launch_count() - l0)
    ms_total = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([ms_total], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    per_step_units = 2 * npoly * L * world
    value = per_step_units * args.steps / (ms_total * 1e-3)
    assert NOCHECK or np.array_equal(d[:L * N].cpu().numpy().view(np.uint64).reshape(L, N), ref0), "data corrupted by timed loop"

    # per-kernel timings (each of the four kernels over the whole batch), CUDA events on the launch stream
    kern = {}
    reps = 5
    for inv, ps, name in [(0, 1, "k_fwd_col"), (0, 2, "k_fwd_row"), (1, 1, "k_inv_row"), (1, 2, "k_inv_col")]:
        for _ in range(2):
            lib.b200fhe_debug_nwt_pass(ctx._h, dp, npoly, L, 0, inv, ps, sp)
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record(stream)
        for _ in range(reps):
            lib.b200fhe_debug_nwt_pass(ctx._h, dp, npoly, L, 0, inv, ps, sp)
        a1.record(stream)
        torch.cuda.synchronize()
        kern[name] = a0.elapsed_time(a1) / reps
    # ABFT overhead: forward transform of the batch, unprotected vs protected (fused checksum), same data
    verdict = torch.zeros(npoly * L, dtype=torch.int32, device="cuda")
    vp_ = verdict.data_ptr()
    def fwd_plain():
        pkg.binding.check(lib.b200fhe_nwt_forward(ctx._h, dp, npoly, L, 0, sp))
    def fwd_prot():
        pkg.binding.check(lib.b200fhe_nwt_forward_protected(ctx._h, dp, npoly, L, 0, vp_, None, None, 0, sp))
    def inv_plain():
        pkg.binding.check(lib.b200fhe_nwt_backward(ctx._h, dp, npoly, L, 0, sp))
    def inv_prot():
        pkg.binding.check(lib.b200fhe_nwt_backward_protected(ctx._h, dp, npoly, L, 0, vp_, None, sp))
    abft_ms = {}
    flagged_inv = 0
    for name, fn in (("inverse", inv_plain), ("inverse_protected", inv_prot), ("forward", fwd_plain), ("forward_protected", fwd_prot)):
        for _ in range(3):
            fn()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record(stream)
        for _ in range(max(args.steps, 5)):
            fn()
        a1.record(stream)
        torch.cuda.synchronize()
        abft_ms[name] = a0.elapsed_time(a1) / max(args.steps, 5)
        if name == "inverse_protected":
            flagged_inv = int(verdict.sum().item())
    abft = {"forward_ms": abft_ms["forward"], "forward_protected_ms": abft_ms["forward_protected"],
            "overhead_pct": 100.0 * (abft_ms["forward_protected"] / abft_ms["forward"] - 1.0),
            "faults_flagged_clean_run": int(verdict.sum().item()),
            "inverse": {"ms": abft_ms["inverse"], "protected_ms": abft_ms["inverse_protected"],
                        "overhead_pct": 100.0 * (abft_ms["inverse_protected"] / abft_ms["inverse"] - 1.0), "faults_flagged_clean_run": flagged_inv},
            "note": "t(protected)/t(unprotected)-1 on the same %d-limb batch; checksum fused into first load / last store" % (npoly * L)}
    peak, peak_src = peaks()
    dom = max(kern, key=kern.get)
    # one pass kernel = half of a limb-transform's 16 B/coeff (the pair shares the 1 MiB)
    algo_bytes_launch = npoly * L * ALGO_BYTES_PER_LIMB_TRANSFORM / 2
    ach = algo_bytes_launch / (kern[dom] * 1e-3) / 1e9
    traffic, traffic_src = newest_ncu_traffic()
    bf_cyc, bf_src = measured_butterfly_cycles()
    sm_hz = 1e6 * (clocks.get("sm_mhz") or 1965.0) if clocks else 1.965e9
    roof = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
            "traffic": traffic.get(dom) if npoly == NPOLY else None,
            "traffic_source": (traffic_src + " (ncu --set full of this command, dram bytes read + written per launch; parsed at run time)") if traffic_src else None,
            "peak_source": peak_src, "algorithmic_bytes_per_launch": algo_bytes_launch,
            "kernel_ms": kern, "note": "one pass kernel over the whole batch = half of a transform's 16 B/coeff; "
            "transform_frac = limb-transforms/s x 1 MiB / peak.  The kernels are bound by the integer-multiply pipe, not by HBM: DESIGN.md section 6",
            "transform_frac": value / world * ALGO_BYTES_PER_LIMB_TRANSFORM / 1e9 / peak,
            "integer_pipe": {"measured_cycles_per_warp_butterfly": bf_cyc, "source": bf_src,
                             "ceiling_limb_transforms_per_s": (592 * sm_hz / (N // 64 * LOG_N * bf_cyc)) if bf_cyc else None,
                             "note": "register-only microbenchmark of the product's butterfly form; N/64 x log2 N warp-butterflies per limb-transform over 592 sub-partitions at the sampled SM clock"}}

    # e2e: host buffers through the C ABI (H2D + kernels + D2H per call)
    hp = hv.ctypes.data_as(C.POINTER(C.c_uint64))
    def e2e_step():
        pkg.binding.check(lib.b200fhe_nwt_forward_host(ctx._h, hp, npoly, L, 0))
        pkg.binding.check(lib.b200fhe_nwt_backward_host(ctx._h, hp, npoly, L, 0))
    for _ in range(2):
        e2e_step()
    barrier()
    e2e_steps = max(3, min(args.steps, 10))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    tt = torch.tensor([dt], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    e2e_val = per_step_units * e2e_steps / float(tt.item())
    assert NOCHECK or np.array_equal(hv[0], ref0), "e2e round trip failed"
    bytes_dir = 2 * npoly * L * N * 8      # fwd call + inv call, each copies the batch in and out

    # the other BASELINE configs, measured in this same process (every rank runs them; rank 0 reports its own)
    sub = {}
    if not args.no_sub:
        import bench_sub
        del d
        torch.cuda.empty_cache()
        sub["pmult"] = bench_sub.pmult_record(pkg, torch, O, peak, local, check=not NOCHECK)
        sub["modup"], sub["keyswitch"] = bench_sub.keyswitch_records(pkg, torch, O, peak, local, check=not NOCHECK)
        torch.cuda.empty_cache()
        sub["fold_check"] = bench_sub.fold_record(pkg, torch, O, peak, local, check=not NOCHECK)
        sub["bgv"] = bench_sub.bgv_record(pkg, torch, peak, local)
        if world > 1:
            torch.cuda.empty_cache()
            sub["modup_sharded"] = bench_sub.modup_sharded_record(pkg, torch, dist, rank, world, local, peak)
            barrier()

    if rank == 0:
        out = {"metric": "NTT limb-transforms/s", "value": value, "unit": "limb-transforms/s", "n_gpus": world,
               "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps,
               "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
               "config": config, "clocks": clocks, "gpu_launches": launches,
               "e2e": {"value": e2e_val, "unit": "limb-transforms/s", "h2d_bytes_per_step": bytes_dir, "d2h_bytes_per_step": bytes_dir,
                       "steps": e2e_steps, "timer": "host perf_counter around synchronous C-ABI calls, max over ranks"},
               "roofline": roof, "abft": abft}
        out.update(sub)
        if world == 1 and not args.no_cpu_baseline:
            v, ms, cores, sample = cpu_reference(None, 1)      # about 10 s of host work
            out["cpu_baseline"] = {"value": v, "unit": "limb-transforms/s", "cores": cores, "kind": "port", "sample": sample}
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
