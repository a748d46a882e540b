"""Sub-records of bench.py's JSON line: the other BASELINE configs measured in the same driver-run
process as the headline, each with its time, its roofline fraction and a parity flag against the oracle.

  pmult      C3  sum of 64 ct x pt, N=2^16, L=36, 2 polys per ciphertext          (HBM-bound)
  modup      C5  key-switch ModUp of one polynomial, N=2^17, 48 + 16 limbs, dnum=3  (integer/tensor bound; 1 GPU: CUDA graph)
  keyswitch  f1  whole hybrid key-switch at C5's shape                             (same)

All device-resident, CUDA events on the launching stream, inputs larger than L2 or rewritten every step.
The oracle (test infrastructure) is used here as the checker only."""
import numpy as np


def _timed(torch, fn, steps, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    st = torch.cuda.current_stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(steps):
        fn()
    e1.record(st)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def _graphed(torch, fn):
    """fn replayed as one CUDA graph, or None when capture is not possible"""
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            fn()
        torch.cuda.current_stream().wait_stream(side)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, capture_error_mode="thread_local"):      # NCCL's helper threads may call CUDA while we capture
            fn()
        return g.replay
    except Exception:
        torch.cuda.synchronize()
        return None


def pmult_record(pkg, torch, O, peak, device, steps=10, check=True):
    log_n, L, T, P = 16, 36, 64, 2
    N = 1 << log_n
    moduli = pkg.primes_create(N, [60] * L)
    ctx = pkg.Context(log_n, moduli, device=device)
    g = torch.Generator(device="cuda").manual_seed(20250606)
    ct = torch.randint(0, 1 << 59, (T, P, L, N), dtype=torch.int64, device="cuda", generator=g)
    pt = torch.randint(0, 1 << 59, (T, L, N), dtype=torch.int64, device="cuda", generator=g)
    out = torch.zeros((P, L, N), dtype=torch.int64, device="cuda")
    cps = [ct[k].data_ptr() for k in range(T)]
    pps = [pt[k].data_ptr() for k in range(T)]
    l0 = pkg.lib().b200fhe_launch_count()
    sp = lambda: torch.cuda.current_stream().cuda_stream
    ctx.pmult_accum(cps, pps, out.data_ptr(), P, L=L, stream=sp())
    torch.cuda.synchronize()
    launches = int(pkg.lib().b200fhe_launch_count() - l0)
    parity = None
    if check:      # every limb, both polynomials, a window of 1024 coefficients per limb (the full-size comparison is tests/test_baseline_shapes_gpu.py)
        w0, w1 = 4096, 4096 + 1024
        cth = np.ascontiguousarray(ct[:, :, :, w0:w1].cpu().numpy().view(np.uint64))
        pth = np.ascontiguousarray(pt[:, :, w0:w1].cpu().numpy().view(np.uint64))
        want = O.pmult_accum(cth, pth, moduli)
        parity = bool(np.array_equal(out[:, :, w0:w1].cpu().numpy().view(np.uint64), want))
    ms = _timed(torch, lambda: ctx.pmult_accum(cps, pps, out.data_ptr(), P, L=L, stream=sp()), steps)
    algo = (3 * T + 2) * L * N * 8
    rec = {"workload": "C3: sum of 64 ct x pt, N=2^16, L=36, 2 polys per ct (dotprod_test.cu shape)", "ms": ms, "launches_per_call": launches,
           "value": T / (ms * 1e-3), "unit": "ct*pt terms/s",
           "roofline": {"bound": "hbm", "achieved": algo / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "frac": algo / (ms * 1e-3) / 1e9 / peak,
                        "algorithmic_bytes_per_call": algo},
           "parity_vs_oracle": parity, "parity_scope": "36 limbs x 2 polys x 1024 coefficients, bit-exact"}
    del ct, pt, out
    return rec


def keyswitch_records(pkg, torch, O, peak, device, steps=10, check=True):
    """(modup record, keyswitch record) at C5's shape on one GPU."""
    log_n, n_q, n_p, dnum = 17, 48, 16, 3
    N, total, alpha = 1 << log_n, n_q + n_p, n_q // dnum
    moduli = pkg.primes_create(N, [60] * total)
    ctx = pkg.Context(log_n, moduli, device=device)
    ks = pkg.KeySwitch(ctx, n_q, n_p, dnum)
    rng = np.random.default_rng(20250606)
    c = np.stack([rng.integers(0, q, size=N, dtype=np.uint64) for q in moduli[:n_q]])
    evk = np.stack([rng.integers(0, q, size=(dnum, 2, N), dtype=np.uint64) for q in moduli]).transpose(1, 2, 0, 3).copy()   # [dnum][2][total][N]
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).cuda()
    d_c, d_eq, d_ep = dev(c), dev(evk[:, :, :n_q]), dev(evk[:, :, n_q:])
    d_out = torch.zeros((2, n_q, N), dtype=torch.int64, device="cuda")
    d_up = torch.zeros((dnum, total, N), dtype=torch.int64, device="cuda")
    sp = lambda: torch.cuda.current_stream().cuda_stream
    lib = pkg.lib()
    l0 = lib.b200fhe_launch_count()
    ks.modup(d_c.data_ptr(), d_up.data_ptr(), stream=sp())
    l1 = lib.b200fhe_launch_count()
    ks.apply(d_c.data_ptr(), d_eq.data_ptr(), d_ep.data_ptr(), d_out.data_ptr(), stream=sp())
    torch.cuda.synchronize()
    l2 = lib.b200fhe_launch_count()
    par_up = par_ks = None
    if check:
        tabs = [O.Tables(log_n, q) for q in moduli]
        got_up = d_up.cpu().numpy().view(np.uint64)
        coeff = O.ntt_inverse_batch(c[:alpha], tabs[:alpha])                       # digit 0, every target limb
        tl = list(range(alpha, total))
        conv = O.bconv_hps(coeff, moduli[:alpha], [moduli[g] for g in tl])
        par_up = bool(np.array_equal(got_up[0][tl], O.ntt_forward_batch(conv, [tabs[g] for g in tl])) and np.array_equal(got_up[0][:alpha], c[:alpha]))
        want = O.keyswitch_hybrid(c, evk, moduli, n_q, n_p, dnum, tabs)
        par_ks = bool(np.array_equal(want, d_out.cpu().numpy().view(np.uint64)))
    f_up = lambda: ks.modup(d_c.data_ptr(), d_up.data_ptr(), stream=sp())
    f_ks = lambda: ks.apply(d_c.data_ptr(), d_eq.data_ptr(), d_ep.data_ptr(), d_out.data_ptr(), stream=sp())
    g_up, g_ks = _graphed(torch, f_up), _graphed(torch, f_ks)
    ms_up = _timed(torch, g_up or f_up, steps)
    ms_ks = _timed(torch, g_ks or f_ks, steps)
    # the checked flavours (eager): conversions only (f2 inside f1), then every limb-transform as well
    ms_eager = _timed(torch, f_ks, steps)
    verdict = torch.zeros((dnum + 2, total), dtype=torch.int32, device="cuda")
    vntt = torch.zeros(ks.ntt_count(), dtype=torch.int32, device="cuda")
    f_prot = lambda: ks.apply_protected(d_c.data_ptr(), d_eq.data_ptr(), d_ep.data_ptr(), d_out.data_ptr(), verdict.data_ptr(), stream=sp())
    f_full = lambda: ks.apply_protected_full(d_c.data_ptr(), d_eq.data_ptr(), d_ep.data_ptr(), d_out.data_ptr(), verdict.data_ptr(), vntt.data_ptr(), stream=sp())
    ref_out = d_out.clone()
    ms_prot = _timed(torch, f_prot, steps)
    same_prot = bool(torch.equal(ref_out, d_out))
    ms_full = _timed(torch, f_full, steps)
    same_full = bool(torch.equal(ref_out, d_out))
    flagged = int(verdict.abs().sum().item()) + int(vntt.abs().sum().item())
    t_up = n_q + dnum * (total - alpha)
    mm_up = dnum * alpha * (1 + total - alpha) * N
    t_ks = t_up + 2 * n_p + 2 * n_q
    mm_ks = mm_up + 2 * n_p * (1 + n_q) * N
    # algorithmic HBM bytes: each limb-transform 16 N, base conversion reads m and writes K limbs per digit, inner product streams the key once
    b_up = (t_up * 16 * N) + dnum * (alpha + total - alpha) * N * 8
    b_ks = b_up + (2 * n_p + 2 * n_q) * 16 * N + 2 * (n_p + n_q) * N * 8 + (dnum * 3 * total + 2 * total) * N * 8
    rec_up = {"workload": "C5: key-switch ModUp of one polynomial (INTT + HPS conversion + NTT), N=2^17, 48 + 16 limbs, dnum=3, 1 GPU",
              "ms": ms_up, "cuda_graph": g_up is not None, "launches_per_call": int(l1 - l0), "value": 1e3 / ms_up, "unit": "polys/s",
              "work_per_call": {"limb_transforms": t_up, "bconv_mulmods": mm_up},
              "roofline": {"bound": "hbm", "achieved": b_up / (ms_up * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "frac": b_up / (ms_up * 1e-3) / 1e9 / peak,
                           "algorithmic_bytes_per_call": b_up, "note": "integer-multiplier / tensor-pipe bound in practice (DESIGN.md 5.5)"},
              "parity_vs_oracle": par_up, "parity_scope": "digit 0: all 64 limbs, bit-exact (all digits: tests/test_baseline_shapes_gpu.py)"}
    rec_ks = {"workload": "f1: hybrid key-switch (ModUp, evk inner product, ModDown), N=2^17, 48 + 16 limbs, dnum=3, 1 GPU",
              "ms": ms_ks, "cuda_graph": g_ks is not None, "launches_per_call": int(l2 - l1), "value": 1e3 / ms_ks, "unit": "key-switches/s",
              "work_per_call": {"limb_transforms": t_ks, "bconv_mulmods": mm_ks},
              "roofline": {"bound": "hbm", "achieved": b_ks / (ms_ks * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "frac": b_ks / (ms_ks * 1e-3) / 1e9 / peak,
                           "algorithmic_bytes_per_call": b_ks},
              "parity_vs_oracle": par_ks, "parity_scope": "both output polynomials, all 48 limbs, bit-exact against the oracle composition",
              "protected": {"ms_eager_unprotected": ms_eager,
                            "conversions_checked": {"ms_eager": ms_prot, "overhead_pct": 100.0 * (ms_prot / ms_eager - 1.0), "checks": int(verdict.numel()),
                                                    "same_output_bits": same_prot},
                            "everything_checked": {"ms_eager": ms_full, "overhead_pct": 100.0 * (ms_full / ms_eager - 1.0),
                                                   "checks": int(verdict.numel() + vntt.numel()), "same_output_bits": same_full},
                            "faults_flagged_clean_run": flagged,
                            "note": "b200fhe_keyswitch_apply_protected / _apply_protected_full: checksum-carrying conversions (tcgen05 contraction with "
                                    "the sums fused), then a fused-checksum verdict per limb-transform as well; the evk inner product is not checked"}}
    ks.close()
    return rec_up, rec_ks


def fold_record(pkg, torch, O, peak, device, steps=10, check=True):
    """f4: element-wise Barrett product with the per-group XOR signatures / the mod 2^k + 1 residue checks, against the
    plain product on the same buffers (2^25 elements = 256 MiB per operand: larger than L2)."""
    import ctypes as C
    q, n_vec, vec_len, fold, k = 68719476767, 1 << 12, 1 << 13, 64, 16      # barrett_final.cpp defaults: 37-bit q, vectors of 8192
    n = n_vec * vec_len
    g = torch.Generator(device="cuda").manual_seed(20250606)
    a = torch.randint(0, q, (n_vec, vec_len), dtype=torch.int64, device="cuda", generator=g)
    b = torch.randint(0, q, (n_vec, vec_len), dtype=torch.int64, device="cuda", generator=g)
    c = torch.zeros_like(a)
    groups = n_vec * (vec_len // fold)
    s1 = torch.zeros(groups, dtype=torch.int64, device="cuda"); s2 = torch.zeros_like(s1)
    v = torch.zeros(groups, dtype=torch.int32, device="cuda")
    sp = lambda: torch.cuda.current_stream().cuda_stream
    lib = pkg.lib()

    def run(mode):
        rc = lib.b200fhe_barrett_fold(q, a.data_ptr(), b.data_ptr(), c.data_ptr(), n_vec, vec_len, fold, k, mode, 0, None, 0,
                                      s1.data_ptr(), s2.data_ptr(), v.data_ptr(), sp())
        assert rc == 0, lib.b200fhe_last_error()
    parity = None
    if check:
        run(1)
        torch.cuda.synchronize()
        w = 64                                                               # the first 64 vectors against the oracle
        want = O.barrett_fold(q, a[:w].cpu().numpy().view(np.uint64), b[:w].cpu().numpy().view(np.uint64), fold, k=k)
        gw = w * (vec_len // fold)
        parity = bool(np.array_equal(c[:w].cpu().numpy().view(np.uint64), want[0]) and np.array_equal(s1[:gw].cpu().numpy().view(np.uint64), want[1])
                      and np.array_equal(s2[:gw].cpu().numpy().view(np.uint64), want[2]) and not bool(v.any().item()))
    ms_plain = _timed(torch, lambda: run(-1), steps)
    ms_sig = _timed(torch, lambda: run(0), steps)
    ms_chk = _timed(torch, lambda: run(1), steps)
    algo = 3 * n * 8
    return {"workload": "f4: element-wise Barrett product (37-bit q, 4096 vectors of 8192) with XOR signatures per group of 64 / with mod 2^16+1 residue checks",
            "ms_plain": ms_plain, "ms_signatures": ms_sig, "ms_residue_checks": ms_chk,
            "overhead_signatures_pct": 100 * (ms_sig / ms_plain - 1), "overhead_residue_checks_pct": 100 * (ms_chk / ms_plain - 1),
            "roofline": {"bound": "hbm", "achieved": algo / (ms_plain * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "frac": algo / (ms_plain * 1e-3) / 1e9 / peak,
                         "algorithmic_bytes_per_call": algo, "note": "plain product; the checked variants move the same bytes + 20 B per group"},
            "parity_vs_oracle": parity, "parity_scope": "64 vectors: products, both signatures, zero verdicts, bit-exact"}


def modup_sharded_record(pkg, torch, dist, rank, world, local, peak, log_n=17, n_q=48, n_p=16, dnum=3, steps=20, warmup=3, graph=True):
    """C5 through the C ABI's limb-sharded ModUp (b200fhe_modup_sharded: one NCCL all-gather per digit on the plan's
    communication stream, overlapped with the previous digit's contraction + NTT).  Every rank first checks its owned
    output limbs bit for bit against the UNSHARDED b200fhe_keyswitch_modup run on its own GPU; timing = CUDA events on
    the launching stream, barrier on both sides, max over ranks.  Returns the record on every rank."""
    B = pkg.binding
    N, total, alpha = 1 << log_n, n_q + n_p, n_q // dnum
    moduli = pkg.primes_create(N, [60] * total)
    rng = np.random.default_rng(20250606)                                  # identical on every rank
    c = np.stack([rng.integers(0, q, size=N, dtype=np.uint64) for q in moduli[:n_q]])
    shard = B.Shard(rank, world, local, dist if world > 1 else None)
    plan = B.ModUpSharded(shard, log_n, moduli, n_q, n_p, dnum)
    own = plan.own_limbs
    own_q = [g for g in own if g < n_q]
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).cuda()
    d_c = dev(c[own_q])
    d_up = torch.zeros((dnum, len(own), N), dtype=torch.int64, device="cuda")
    sp = lambda: torch.cuda.current_stream().cuda_stream
    f = lambda: plan.run(d_c.data_ptr(), d_up.data_ptr(), stream=sp())
    lib = pkg.lib()
    l0 = lib.b200fhe_launch_count()
    f()
    torch.cuda.synchronize()
    launches = int(lib.b200fhe_launch_count() - l0)
    # parity gate: the unsharded key-switch ModUp on this rank's GPU (itself checked against the oracle in bench_sub.keyswitch_records / the tests)
    ctx = pkg.Context(log_n, moduli, device=local)
    ks = pkg.KeySwitch(ctx, n_q, n_p, dnum)
    d_all, d_ref = dev(c), torch.zeros((dnum, total, N), dtype=torch.int64, device="cuda")
    ks.modup(d_all.data_ptr(), d_ref.data_ptr(), stream=sp())
    torch.cuda.synchronize()
    ok = torch.tensor([1 if torch.equal(d_up, d_ref[:, own]) else 0], device="cuda")
    if world > 1:
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    parity = bool(int(ok.item()))
    ks.close(); del d_all, d_ref, ctx
    torch.cuda.empty_cache()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    step, graphed = f, False
    # world > 1 stays eager: capturing the call with the NCCL all-gather inside never returned on this stack (torch 2.11 /
    # NCCL 2.28.9, both ranks capturing the same sequence; killed by the 300 s timeout in run r2g), so it is not attempted
    if graph and world == 1:
        g = _graphed(torch, f)
        if g is not None:
            step, graphed = g, True
    for _ in range(max(warmup, 3)):
        step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st = torch.cuda.current_stream()
    e0.record(st)
    for _ in range(steps):
        step()
    e1.record(st)
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / steps
    k_out = total - alpha
    rec = {"workload": "C5: ModUp of one polynomial, N=2^%d, %d + %d limbs, dnum=%d, limb g on rank g %% %d, through b200fhe_modup_sharded (C ABI)"
                       % (log_n, n_q, n_p, dnum, world),
           "ms": ms, "value": 1e3 / ms, "unit": "polys/s", "n_gpus": world, "scaling": "strong", "cuda_graph": graphed,
           "launches_per_call_per_rank": launches,
           "collective": ("ncclAllGather of %d scaled limbs per digit (%d MiB), %d per call, on the plan's communication stream" % (alpha, alpha * N * 8 >> 20, dnum))
                         if world > 1 else "none",
           "work_per_call": {"limb_transforms": n_q + dnum * k_out, "bconv_mulmods": dnum * alpha * (1 + k_out) * N,
                             "allgather_bytes": dnum * alpha * N * 8 if world > 1 else 0},
           "parity_vs_unsharded": parity, "parity_scope": "every owned limb of every digit on every rank, bit-exact against b200fhe_keyswitch_modup"}
    plan.close(); shard.close()
    return rec


def bgv_record(pkg, torch, peak, device, steps=10):
    """f3: the homomorphic steps of dotprod_test.cu:113-152 at its own parameters (N=16384, six 50-bit primes of which two special,
    20-bit batching plain modulus): multiply, relinearize, mod-switch, rotate, timed per call on device-resident ciphertexts; the
    pipeline's result is decrypted and compared with the slot-wise product (the reference's own correctness criterion)."""
    log_n, N = 14, 1 << 14
    tq = pkg.primes_create(N, [50] * 6)
    tt = pkg.primes_create(N, [20])[0]
    bgv = pkg.binding.Bgv(log_n, tq, 2, tt, seed=7, device=device)
    rng = np.random.default_rng(3)
    xa, xb = rng.integers(0, 100, N).astype(np.uint64), rng.integers(0, 100, N).astype(np.uint64)
    ca = torch.zeros((2, 4, N), dtype=torch.int64, device="cuda"); cb = torch.zeros_like(ca)
    bgv.encrypt(bgv.encode(xa), ca.data_ptr()); bgv.encrypt(bgv.encode(xb), cb.data_ptr())
    c3 = torch.zeros((3, 4, N), dtype=torch.int64, device="cuda")
    lo = torch.zeros((2, 3, N), dtype=torch.int64, device="cuda")
    lib = pkg.lib()
    l0 = lib.b200fhe_launch_count()
    bgv.multiply(ca.data_ptr(), cb.data_ptr(), c3.data_ptr(), 4)
    bgv.relinearize(c3.data_ptr(), 4)
    cf = bgv.mod_switch_to_next(c3[:2].contiguous().data_ptr(), 2, 4, lo.data_ptr())
    bgv.rotate(lo.data_ptr(), 3, 1)
    torch.cuda.synchronize()
    l1 = lib.b200fhe_launch_count()
    prod = xa * xb % np.uint64(tt)
    want = np.concatenate([np.roll(prod[:N // 2], -1), np.roll(prod[N // 2:], -1)])
    ok = bool((bgv.decode(bgv.decrypt(lo.data_ptr(), 2, 3, cf)) == want).all())
    two = c3[:2].contiguous()
    scratch = torch.zeros_like(c3)
    def relin():
        scratch.copy_(c3_fresh)
        bgv.relinearize(scratch.data_ptr(), 4)
    bgv.multiply(ca.data_ptr(), cb.data_ptr(), c3.data_ptr(), 4)
    c3_fresh = c3.clone()
    ms = {"multiply": _timed(torch, lambda: bgv.multiply(ca.data_ptr(), cb.data_ptr(), c3.data_ptr(), 4), steps),
          "copy_3x4_limbs": _timed(torch, lambda: scratch.copy_(c3_fresh), steps),
          "copy_plus_relinearize": _timed(torch, relin, steps),
          "mod_switch_to_next": _timed(torch, lambda: bgv.mod_switch_to_next(two.data_ptr(), 2, 4, lo.data_ptr()), steps),
          "rotate_rows_1": _timed(torch, lambda: bgv.rotate(lo.data_ptr(), 3, 1), steps)}
    bgv.close()
    return {"workload": "f3: BGV steps of dotprod_test.cu (N=16384, {50 x 6} with 2 special primes, t = 20-bit batching prime), device-resident",
            "ms": ms, "launches_multiply_relin_modswitch_rotate": int(l1 - l0), "decrypts_to_slotwise_product_rotated": ok,
            "note": "per call through b200fhe_bgv_* (what shim/phantom_bgv.h forwards to); relinearize works in place, so it is timed with the "
                    "copy that restores its input (the copy alone is listed)"}

