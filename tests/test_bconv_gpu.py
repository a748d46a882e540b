"""GPU parity: ModUp base conversion (config C5) through the C ABI vs the oracle, bit-exact, and the
single-GPU form of the limb-sharded ModUp pipeline (INTT -> scale -> [gather] -> contract -> NTT)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("log_n,bits,m,K", [
    (4, 30, 2, 3), (10, 50, 3, 5), (12, 60, 4, 4), (12, 60, 16, 7), (13, 50, 6, 2), (12, 60, 20, 9), (14, 60, 16, 48),
])
def test_modup_host_vs_oracle(pkg, O, log_n, bits, m, K):
    N = 1 << log_n
    moduli = pkg.primes_create(N, [bits] * (m + K))
    ctx = pkg.Context(log_n, moduli)
    src, dst = list(range(m)), list(range(m, m + K))
    bc = pkg.BaseConv(ctx, src, dst)
    x = O.uniform_limbs(moduli[:m], N, seed=log_n * 100 + m)
    got = bc.modup_host(x)
    want = O.bconv_hps(x, moduli[:m], moduli[m:])
    assert np.array_equal(got, want)


def test_modup_interleaved_indices_and_edge_values(pkg, O):
    """source / target moduli interleaved in the context; inputs 0, p-1 and words >= p (taken mod p)."""
    log_n, N = 12, 4096
    moduli = pkg.primes_create(N, [60] * 9)
    ctx = pkg.Context(log_n, moduli)
    src, dst = [7, 1, 4], [0, 8, 2, 3, 5, 6]
    bc = pkg.BaseConv(ctx, src, dst)
    p = [moduli[i] for i in src]; q = [moduli[i] for i in dst]
    x = O.uniform_limbs(p, N, seed=1)
    x[:, 0] = 0
    x[:, 1] = np.array([pi - 1 for pi in p], dtype=np.uint64)
    x[:, 2] = np.uint64(2 ** 64 - 1)
    assert np.array_equal(bc.modup_host(x), O.bconv_hps(x, p, q))


def test_hps_result_is_x_plus_small_multiple_of_P(pkg, O):
    """Property at C5's shape (N=2^17, alpha=16): out_k == (x + e P) mod q_k with 0 <= e < alpha, e the same for all k."""
    log_n, N, m, K = 17, 1 << 17, 16, 4
    moduli = pkg.primes_create(N, [60] * (m + K))
    ctx = pkg.Context(log_n, moduli)
    bc = pkg.BaseConv(ctx, list(range(m)), list(range(m, m + K)))
    x = O.uniform_limbs(moduli[:m], N, seed=5)
    got = bc.modup_host(x)
    P = 1
    for pj in moduli[:m]:
        P *= pj
    # exact CRT of a sample of coefficients with Python integers
    for i in list(range(0, N, N // 16)) + [N - 1]:
        xi = 0
        for j, pj in enumerate(moduli[:m]):
            Pj = P // pj
            xi = (xi + int(x[j, i]) * pow(Pj, -1, pj) % pj * Pj) % P
        es = set()
        for k, qk in enumerate(moduli[m:]):
            e = [e for e in range(m) if (xi + e * P) % qk == int(got[k, i])]
            assert len(e) == 1
            es.add(e[0])
        assert len(es) == 1


@pytest.mark.parametrize("log_n,bits,m,K", [(4, 30, 2, 3), (10, 50, 3, 5), (12, 60, 16, 7), (12, 60, 1, 2), (11, 60, 40, 3)])
def test_exact_crt_mode_vs_oracle(pkg, O, log_n, bits, m, K):
    """b200fhe_bconv_create_mode(..., EXACT): (x mod P) mod q_k as motivation/baseConv.py:67-83, including the values whose
    overshoot count e a floating-point estimate cannot decide (x = 0, 1, 2, P-1, P-2) and the largest overshoot."""
    N = 1 << log_n
    moduli = pkg.primes_create(N, [bits] * (m + K))
    ctx = pkg.Context(log_n, moduli)
    p, q = moduli[:m], moduli[m:]
    P = 1
    for v in p:
        P *= v
    rng = np.random.default_rng(m * 100 + K)
    vals = [int.from_bytes(rng.bytes(8 * m + 8), "little") % P for _ in range(N)]
    vals[:5] = [0, 1, 2 % P, P - 1, (P - 2) % P]
    x = np.array([[v % pj for v in vals] for pj in p], dtype=np.uint64)
    want = np.array([[v % qk for v in vals] for qk in q], dtype=np.uint64)
    assert np.array_equal(O.bconv_exact(x[:, :64], p, q), want[:, :64])          # the restatement agrees with plain arithmetic
    exact = pkg.BaseConv(ctx, list(range(m)), list(range(m, m + K)), mode=pkg.BaseConv.EXACT)
    assert np.array_equal(exact.modup_host(x), want)
    fast = pkg.BaseConv(ctx, list(range(m)), list(range(m, m + K)))
    hps = fast.modup_host(x)
    assert np.array_equal(hps, O.bconv_hps(x, p, q))
    if m > 1:
        assert not np.array_equal(hps, want)                                      # the fast flavour does overshoot somewhere


def test_sharded_pipeline_world1_matches_oracle(pkg, O):
    """modup_digit with world = 1 (no collective): INTT -> HPS conversion -> NTT of every digit == oracle composition."""
    import torch
    from fhe_reliability_gpu_b200 import modup
    log_n, N, n_q, n_p, dnum = 12, 4096, 6, 2, 3
    moduli = pkg.primes_create(N, [60] * (n_q + n_p))
    tabs = [O.Tables(log_n, q) for q in moduli]
    plan = modup.ShardPlan(n_q, n_p, dnum, 1, 0)
    be = modup.CudaBackend(pkg, log_n, moduli, 0, 1, 0)
    a = O.uniform_limbs(moduli[:n_q], N, seed=9)                 # NTT-domain limbs of Q
    mine = {g: torch.from_numpy(a[g].view(np.int64)).cuda() for g in range(n_q)}
    for d in range(dnum):
        res = modup.modup_digit(plan, be, None, d, mine)
        torch.cuda.synchronize()
        dl, tl = plan.digit_limbs(d), plan.target_limbs(d)
        coeff = np.stack([O.ntt_inverse(a[g], tabs[g]) for g in dl])
        conv = O.bconv_hps(coeff, [moduli[g] for g in dl], [moduli[g] for g in tl])
        for i, g in enumerate(tl):
            assert np.array_equal(res[g].cpu().numpy().view(np.uint64), O.ntt_forward(conv[i], tabs[g])), (d, g)
        for g in dl:
            assert np.array_equal(res[g].cpu().numpy().view(np.uint64), a[g])


def test_bconv_rejects_bad_arguments(pkg):
    ctx = pkg.Context(12, pkg.primes_create(4096, [60] * 4))
    with pytest.raises(pkg.B200FHEError):
        pkg.BaseConv(ctx, [0, 0], [1])          # repeated source modulus
    with pytest.raises(pkg.B200FHEError):
        pkg.BaseConv(ctx, [0, 1], [9])          # target outside the context
    with pytest.raises(pkg.B200FHEError):
        pkg.BaseConv(ctx, [], [1])

