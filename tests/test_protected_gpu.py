"""GPU parity: protected forward NTT with the fused ABFT checksum (config C4) vs the oracle.

The checksum is the weighted identity of rfhe_framewk/src/negaclic_ntt.py:130-149 with the
prototype's own weights; faults are injected per operation as in
rfhe_framewk/src/ntt_test/relia_ntt_sim.cpp:169-206 (kinds SBF/DBF/MLF, :109-132).  For a given
plan everything is deterministic, so outputs, both checksums and the verdict must equal the
oracle's bit for bit; the detection statistics are then compared with the simulator's."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ctx(pkg, log_n, moduli):
    return pkg.Context(log_n, moduli)


@pytest.mark.parametrize("log_n,bits,L,n_poly", [
    (4, 20, 2, 1), (6, 30, 1, 3), (8, 60, 2, 1), (11, 50, 2, 2),
    (12, 60, 2, 1), (12, 50, 32, 1), (13, 60, 3, 2), (14, 50, 6, 1), (15, 60, 2, 1), (16, 60, 4, 2), (17, 60, 2, 1),
])
def test_clean_run_matches_oracle_and_reports_no_fault(pkg, O, log_n, bits, L, n_poly):
    N = 1 << log_n
    moduli = pkg.primes_create(N, [bits] * L)
    ctx = _ctx(pkg, log_n, moduli)
    tabs = [O.Tables(log_n, q) for q in moduli]
    a = np.stack([O.uniform_limbs(moduli, N, seed=50 + log_n + p) for p in range(n_poly)])
    got, verdict, sums = ctx.nwt_forward_protected_host(a.copy())
    assert not verdict.any()
    for p in range(n_poly):
        for l in range(L):
            out, det, lhs, rhs = O.ntt_forward_protected(a[p, l], tabs[l])
            assert not det
            assert np.array_equal(got[p, l], out)
            assert (int(sums[p * L + l, 0]), int(sums[p * L + l, 1])) == (lhs, rhs)
    # the unprotected entry point gives the same transform
    assert np.array_equal(ctx.nwt_forward_host(a.copy()), got)


def test_weights_match_the_prototype(pkg, O):
    for log_n, bits in [(6, 30), (8, 60), (12, 60), (13, 50)]:
        N = 1 << log_n
        q = pkg.primes_create(N, [bits])[0]
        ctx = _ctx(pkg, log_n, [q])
        w, wh = ctx.abft_weights(0)
        ow, oom = O.abft_weights(O.Tables(log_n, q))
        assert np.array_equal(w, ow) and np.array_equal(wh, oom)


@pytest.mark.parametrize("log_n,bits", [(6, 30), (10, 60), (12, 60), (13, 50), (14, 60), (16, 60), (17, 60)])
def test_injected_faults_match_oracle_bit_for_bit(pkg, O, log_n, bits):
    """Every stage is hit at least once; outputs, checksums and verdicts equal the oracle's."""
    N = 1 << log_n
    L = 2
    moduli = pkg.primes_create(N, [bits] * L)
    ctx = _ctx(pkg, log_n, moduli)
    tabs = [O.Tables(log_n, q) for q in moduli]
    weights = [O.abft_weights(t) for t in tabs]
    rng = np.random.default_rng(log_n)
    a = O.uniform_limbs(moduli, N, seed=77)
    stages = list(range(log_n))
    for rep in range(3):
        rng.shuffle(stages)
        for s0 in range(0, log_n, 4):
            plan_gpu, plan_cpu = [], {0: [], 1: []}
            for st in stages[s0:s0 + 4]:
                limb = int(rng.integers(0, L))
                bf = int(rng.integers(0, N // 2))
                kind = int(rng.integers(1, 3))
                payload = int(rng.integers(0, 2 ** 63)) if kind == 2 else (1 << int(rng.integers(0, bits)))
                plan_gpu.append((limb, st, bf, kind, payload))
                plan_cpu[limb].append((st, bf, kind, payload))
            got, verdict, sums = ctx.nwt_forward_protected_host(a.copy(), plan=plan_gpu)
            for l in range(L):
                out, det, lhs, rhs = O.ntt_forward_protected(a[l], tabs[l], plan_cpu[l], weights[l])
                assert np.array_equal(got[l], out), (log_n, plan_gpu, l)
                assert (int(sums[l, 0]), int(sums[l, 1])) == (lhs, rhs)
                assert bool(verdict[l]) == det


def test_detection_statistics_match_relia_ntt_sim(pkg, O):
    """relia_ntt_sim --N 64 --q-bits 10 (Q=769): single-fault miss rates are ~1/Q (reliaNTT_test_result.txt:7-9:
    0.001304 / 0 / 0.001275).  Same experiment here: one SBF / DBF / MLF per trial at a uniformly random operation.
    Per trial the GPU verdict must equal the oracle's; the miss rate must be consistent with 1/Q."""
    log_n, N, Q = 6, 64, 769
    assert O.lib().orc_find_ntt_prime(N, 10) == Q
    ctx = _ctx(pkg, log_n, [Q])
    tab = O.Tables(log_n, Q)
    weights = O.abft_weights(tab)
    rng = np.random.default_rng(2025)
    trials, per_call = 6000, 16
    miss = {1: 0, 2: 0, 3: 0}
    count = {1: 0, 2: 0, 3: 0}
    effective = 0
    for t0 in range(0, trials, per_call):
        a = np.stack([rng.integers(0, Q, size=(1, N), dtype=np.uint64) for _ in range(per_call)])
        plan_gpu, plans = [], []
        for i in range(per_call):
            fk = 1 + (t0 + i) % 3                      # SBF, DBF, MLF
            st, bf = int(rng.integers(0, log_n)), int(rng.integers(0, N // 2))
            if fk == 1:
                kind, payload = 1, 1 << int(rng.integers(0, 10))
            elif fk == 2:
                b1, b2 = rng.choice(10, size=2, replace=False)
                kind, payload = 1, (1 << int(b1)) | (1 << int(b2))
            else:
                kind, payload = 2, int(rng.integers(0, 2 ** 63))
            plan_gpu.append((i, st, bf, kind, payload))
            plans.append((fk, (st, bf, kind, payload)))
        got, verdict, sums = ctx.nwt_forward_protected_host(a.copy(), plan=plan_gpu)
        for i, (fk, pl) in enumerate(plans):
            out, det, lhs, rhs = O.ntt_forward_protected(a[i, 0], tab, [pl], weights)
            assert np.array_equal(got[i, 0], out) and bool(verdict[i]) == det
            clean = O.ntt_forward(a[i, 0], tab)
            if not np.array_equal(out, clean):        # the fault changed the result (v' != v)
                effective += 1
                count[fk] += 1
                miss[fk] += (not det)
    total_miss = sum(miss.values())
    rate = total_miss / effective
    # binomial around 1/Q: mean ~ effective/769 (~7.5), allow a generous window; never "always detected by luck" only
    assert effective > 0.95 * trials
    assert rate < 5.0 / Q, (miss, count)


def test_harness_style_input_flip_is_not_flagged(pkg, O):
    """ntt_test.cu:105-135 corrupts the INPUT buffer before the transform: ABFT of the transform cannot (and must
    not) flag that -- the checksum is taken over what the kernel loads.  Documented in DESIGN.md."""
    log_n, N = 12, 4096
    moduli = pkg.primes_create(N, [50] * 2)
    ctx = _ctx(pkg, log_n, moduli)
    a = O.uniform_limbs(moduli, N, seed=3)
    a[0, 17] ^= np.uint64(1 << 63)
    got, verdict, _ = ctx.nwt_forward_protected_host(a.copy())
    assert not verdict.any()
    tabs = [O.Tables(log_n, q) for q in moduli]
    assert np.array_equal(got, O.ntt_forward_batch(a % np.array(moduli, dtype=np.uint64)[:, None], tabs))


def test_storage_fault_between_the_passes_is_detected_and_stays_in_its_row(pkg, O):
    """b200fhe_inject kind 3: the intermediate the column pass stored to HBM is XOR-ed before the row pass loads it (the SCF class of
    rfhe_framewk/src/four_step_ntt_protected.cpp:163-196 at the 2D transform's round trip).  The checksum trips for that limb
    only, the damage stays inside the 256 outputs of the row the word belongs to, and the input-side checksum is unchanged."""
    from fhe_reliability_gpu_b200.binding import fault_payload
    log_n, L, n_poly = 14, 2, 2
    N = 1 << log_n
    moduli = pkg.primes_create(N, [60] * L)
    ctx = pkg.Context(log_n, moduli)
    rng = np.random.default_rng(77)
    a = np.stack([np.stack([rng.integers(0, q, N, dtype=np.uint64) for q in moduli]) for _ in range(n_poly)])
    clean, v0, s0 = ctx.nwt_forward_protected_host(a.copy())
    assert not v0.any()
    for trial, subtype in enumerate(["SCF-BF", "SCF-MBU", "MCF-PPE", "MCF-CTE"] * 3):
        limb, idx = int(rng.integers(0, n_poly * L)), int(rng.integers(0, N))
        mask = fault_payload(subtype, 60, rng)
        out, v, s = ctx.nwt_forward_protected_host(a.copy(), plan=[(limb, 0, idx, 3, mask)])
        want_v = [0] * (n_poly * L); want_v[limb] = 1
        assert list(v) == want_v, (subtype, limb, idx, hex(mask))
        assert (s[:, 0] == s0[:, 0]).all() and s[limb, 1] != s0[limb, 1]
        diff = (out.reshape(n_poly * L, N) != clean.reshape(n_poly * L, N))
        row = idx // 256
        assert diff[limb, row * 256:(row + 1) * 256].any()
        diff[limb, row * 256:(row + 1) * 256] = False
        assert not diff.any()
    # a storage fault together with a butterfly fault in another limb; and the argument checks
    out, v, _ = ctx.nwt_forward_protected_host(a.copy(), plan=[(0, 0, 17, 3, 1 << 3), (3, 5, 100, 1, 1 << 20)])
    assert list(v) == [1, 0, 0, 1]
    with pytest.raises(pkg.B200FHEError):
        ctx.nwt_forward_protected_host(a.copy(), plan=[(0, 0, N, 3, 1)])
    small = pkg.Context(6, pkg.primes_create(64, [30]))
    with pytest.raises(pkg.B200FHEError):
        small.nwt_forward_protected_host(np.zeros((1, 1, 64), dtype=np.uint64), plan=[(0, 0, 1, 3, 1)])
