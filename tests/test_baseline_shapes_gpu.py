"""GPU parity at the BASELINE.json configs themselves, every limb against the oracle (VERDICT r1 "weak" #1).

C2: N=2^16, L=36 x 60-bit, forward + inverse, several polynomials (the shape bench.py times).
C3: ct x pt accumulate, T=64 terms, N=2^16, L=36, both polynomials of the ciphertext (dotprod_test.cu shape).
C5: ModUp base conversion of one digit, N=2^17, 16 -> 48 limbs (key-switch shape), and the whole C5 ModUp.
Sizes are the full ones; the oracle is OpenMP C and finishes each case in seconds."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_c2_every_limb_of_three_polys_vs_oracle(pkg, O):
    log_n, L, n_poly = 16, 36, 3          # 3 polys: an odd batch exercises the pair and the single path of the row passes
    N = 1 << log_n
    moduli = pkg.primes_create(N, [60] * L)
    ctx = pkg.Context(log_n, moduli)
    bt = O.BatchTables([O.Tables(log_n, q) for q in moduli])
    a = np.stack([O.uniform_limbs(moduli, N, seed=4000 + i) for i in range(n_poly)])
    want = O.ntt_forward_batch(a, bt)
    got = ctx.nwt_forward_host(a.copy())
    assert np.array_equal(got, want)                                   # all 108 limbs, bit for bit
    assert np.array_equal(ctx.nwt_backward_host(got.copy()), a)
    b = np.stack([O.uniform_limbs(moduli, N, seed=5000 + i) for i in range(n_poly)])
    assert np.array_equal(ctx.nwt_backward_host(b.copy()), O.ntt_inverse_batch(b, bt))
    # the bench batch (8 polys) through the device-resident entry points
    import torch
    a8 = np.stack([O.uniform_limbs(moduli, N, seed=6000 + i) for i in range(8)])
    d = torch.from_numpy(a8.view(np.int64)).cuda()
    s = torch.cuda.current_stream().cuda_stream
    lib = pkg.lib()
    pkg.binding.check(lib.b200fhe_nwt_forward(ctx._h, d.data_ptr(), 8, L, 0, s))
    torch.cuda.synchronize()
    assert np.array_equal(d.cpu().numpy().view(np.uint64), O.ntt_forward_batch(a8, bt))
    pkg.binding.check(lib.b200fhe_nwt_backward(ctx._h, d.data_ptr(), 8, L, 0, s))
    torch.cuda.synchronize()
    assert np.array_equal(d.cpu().numpy().view(np.uint64), a8)


def test_c2_protected_every_limb_vs_oracle(pkg, O):
    """C4 at C2's shape: outputs of the protected transform == unprotected oracle, no verdict raised, both checksums equal."""
    log_n, L, n_poly = 16, 36, 2
    N = 1 << log_n
    moduli = pkg.primes_create(N, [60] * L)
    ctx = pkg.Context(log_n, moduli)
    bt = O.BatchTables([O.Tables(log_n, q) for q in moduli])
    a = np.stack([O.uniform_limbs(moduli, N, seed=7000 + i) for i in range(n_poly)])
    got, verdict, sums = ctx.nwt_forward_protected_host(a.copy())
    assert np.array_equal(got, O.ntt_forward_batch(a, bt))
    assert not verdict.any() and (sums[:, 0] == sums[:, 1]).all()
    # input-side checksum against exact arithmetic on a few limbs
    for l in (0, 19, 35):
        w, _ = O.abft_weights(bt.tabs[l])
        want = int(np.sum(w.astype(object) * a[1, l].astype(object)) % moduli[l])
        assert int(sums[L + l, 0]) == want


def test_c3_t64_every_limb_both_polys_vs_oracle(pkg, O):
    log_n, N, L, T, P = 16, 1 << 16, 36, 64, 2
    moduli = pkg.primes_create(N, [60] * L)
    ctx = pkg.Context(log_n, moduli)
    rng = np.random.default_rng(31)
    ct = np.empty((T, P, L, N), dtype=np.uint64); pt = np.empty((T, L, N), dtype=np.uint64)
    for l, q in enumerate(moduli):
        ct[:, :, l, :] = rng.integers(0, q, size=(T, P, N), dtype=np.uint64)
        pt[:, l, :] = rng.integers(0, q, size=(T, N), dtype=np.uint64)
    got = ctx.pmult_accum_host(ct, pt)
    for l0 in range(0, L, 6):                      # oracle over chunks of limbs (memory: contiguous copies of 1/6 of the inputs)
        want = O.pmult_accum(ct[:, :, l0:l0 + 6, :], pt[:, l0:l0 + 6, :], moduli[l0:l0 + 6])
        assert np.array_equal(got[:, l0:l0 + 6, :], want), "limbs %d..%d" % (l0, l0 + 5)


def test_c5_digit_conversion_full_shape_vs_oracle(pkg, O):
    log_n, m, K = 17, 16, 48
    N = 1 << log_n
    moduli = pkg.primes_create(N, [60] * (m + K))
    ctx = pkg.Context(log_n, moduli)
    bc = pkg.BaseConv(ctx, list(range(m)), list(range(m, m + K)))
    x = O.uniform_limbs(moduli[:m], N, seed=1717)
    assert np.array_equal(bc.modup_host(x), O.bconv_hps(x, moduli[:m], moduli[m:]))


def test_c5_whole_modup_full_shape_vs_oracle(pkg, O):
    """Key-switch ModUp at C5's shape (N=2^17, 48 + 16 limbs, dnum=3): INTT -> conversion -> NTT of every digit, every limb."""
    import torch
    log_n, n_q, n_p, dnum = 17, 48, 16, 3
    N = 1 << log_n
    moduli = pkg.primes_create(N, [60] * (n_q + n_p))
    ctx = pkg.Context(log_n, moduli)
    tabs = [O.Tables(log_n, q) for q in moduli]
    c = O.uniform_limbs(moduli[:n_q], N, seed=555)
    ks = pkg.KeySwitch(ctx, n_q, n_p, dnum)
    dc = torch.from_numpy(c.view(np.int64)).cuda()
    up = torch.zeros((dnum, n_q + n_p, N), dtype=torch.int64, device="cuda")
    ks.modup(dc.data_ptr(), up.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    got = up.cpu().numpy().view(np.uint64)
    ks.close()
    alpha = n_q // dnum
    coeff = O.ntt_inverse_batch(c, tabs[:n_q])
    for d in range(dnum):
        dl = list(range(d * alpha, (d + 1) * alpha))
        tl = [g for g in range(n_q + n_p) if g not in dl]
        conv = O.bconv_hps(coeff[dl], [moduli[g] for g in dl], [moduli[g] for g in tl])
        want = O.ntt_forward_batch(conv, [tabs[g] for g in tl])
        assert np.array_equal(got[d][tl], want), "digit %d targets" % d
        assert np.array_equal(got[d][dl], c[dl]), "digit %d own limbs" % d
