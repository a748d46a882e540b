"""GPU parity of the hybrid key-switch (INTT -> ModUp -> NTT -> sum_d up_d (.) evk_d -> INTT ->
ModDown -> NTT; what relinearize_inplace / rotate_inplace run, reliability_test/dotprod_test.cu:114,146):
bit-exact against the oracle's composition of the pinned primitives, and the size-independent property that
makes it a key-switch: with a real RLWE switching key, out0 + out1 s - c s' is a SMALL polynomial."""
import numpy as np
import pytest

from ks_common import check_switch_noise, make_switch_key

pytestmark = pytest.mark.gpu


def _setup(pkg, O, log_n, n_q, n_p, bits=60):
    N = 1 << log_n
    moduli = pkg.primes_create(N, [bits] * (n_q + n_p))
    ctx = pkg.Context(log_n, moduli)
    tabs = [O.Tables(log_n, q) for q in moduli]
    return N, moduli, ctx, tabs


def _dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).cuda()


def _host(t):
    return t.cpu().numpy().view(np.uint64)


def _split(evk, n_q):
    return np.ascontiguousarray(evk[:, :, :n_q]), np.ascontiguousarray(evk[:, :, n_q:])


@pytest.mark.parametrize("log_n,n_q,n_p,dnum", [(10, 6, 2, 3), (12, 4, 4, 1), (12, 6, 1, 6), (13, 8, 3, 2), (11, 9, 3, 3)])
def test_keyswitch_pieces_and_whole_vs_oracle(pkg, O, log_n, n_q, n_p, dnum):
    import torch
    N, moduli, ctx, tabs = _setup(pkg, O, log_n, n_q, n_p)
    total = n_q + n_p
    ks = pkg.KeySwitch(ctx, n_q, n_p, dnum)
    c = O.uniform_limbs(moduli[:n_q], N, seed=log_n + dnum)
    evk = np.stack([np.stack([O.uniform_limbs(moduli, N, seed=1000 + 10 * d + p) for p in range(2)]) for d in range(dnum)])
    # ModUp
    d_c = _dev(torch, c)
    d_up = torch.zeros((dnum, total, N), dtype=torch.int64, device="cuda")
    ks.modup(d_c.data_ptr(), d_up.data_ptr())
    torch.cuda.synchronize()
    want_up = O.keyswitch_modup(c, moduli, n_q, n_p, dnum, tabs)
    assert np.array_equal(_host(d_up), want_up)
    assert np.array_equal(_host(d_c), c)                       # the input is left untouched
    # ModDown of an arbitrary polynomial over Q u P
    acc = O.uniform_limbs(moduli, N, seed=77)
    d_q, d_p = _dev(torch, acc[:n_q]), _dev(torch, acc[n_q:])
    d_o = torch.zeros((n_q, N), dtype=torch.int64, device="cuda")
    ks.moddown(d_q.data_ptr(), d_p.data_ptr(), d_o.data_ptr(), 1)
    torch.cuda.synchronize()
    assert np.array_equal(_host(d_o), O.keyswitch_moddown(acc, moduli, n_q, n_p, tabs))
    # the whole switch
    eq, ep = _split(evk, n_q)
    d_eq, d_ep = _dev(torch, eq), _dev(torch, ep)
    d_out = torch.zeros((2, n_q, N), dtype=torch.int64, device="cuda")
    ks.apply(d_c.data_ptr(), d_eq.data_ptr(), d_ep.data_ptr(), d_out.data_ptr())
    torch.cuda.synchronize()
    assert np.array_equal(_host(d_out), O.keyswitch_hybrid(c, evk, moduli, n_q, n_p, dnum, tabs))
    # replays give the same bits (workspace reuse leaves no state behind)
    d_out2 = torch.zeros_like(d_out)
    ks.apply(d_c.data_ptr(), d_eq.data_ptr(), d_ep.data_ptr(), d_out2.data_ptr())
    torch.cuda.synchronize()
    assert torch.equal(d_out, d_out2)


@pytest.mark.parametrize("log_n,n_q,n_p,dnum", [(11, 6, 2, 3), (12, 8, 4, 2), (15, 12, 4, 3), (17, 48, 16, 3)])
def test_keyswitch_semantics_with_rlwe_key(pkg, O, log_n, n_q, n_p, dnum):
    """out0 + out1 s == c s' + noise, noise the same small integer polynomial modulo every q_l."""
    import torch
    N, moduli, ctx, tabs = _setup(pkg, O, log_n, n_q, n_p)
    rng = np.random.default_rng(log_n * 31 + dnum)
    s = rng.integers(-1, 2, N)
    s_new = rng.integers(-1, 2, N)
    evk = make_switch_key(O, rng, s, s_new, moduli, n_q, n_p, dnum, tabs)
    c = O.uniform_limbs(moduli[:n_q], N, seed=9)
    ks = pkg.KeySwitch(ctx, n_q, n_p, dnum)
    eq, ep = _split(evk, n_q)
    d_c, d_eq, d_ep = _dev(torch, c), _dev(torch, eq), _dev(torch, ep)
    d_out = torch.zeros((2, n_q, N), dtype=torch.int64, device="cuda")
    ks.apply(d_c.data_ptr(), d_eq.data_ptr(), d_ep.data_ptr(), d_out.data_ptr())
    torch.cuda.synchronize()
    out = _host(d_out)
    check_switch_noise(O, out, c, s, s_new, moduli[:n_q], tabs[:n_q])


def test_keyswitch_argument_errors(pkg):
    moduli = pkg.primes_create(1 << 10, [50] * 6)
    ctx = pkg.Context(10, moduli)
    with pytest.raises(pkg.B200FHEError):
        pkg.KeySwitch(ctx, 4, 1, 2)          # n_q + n_p != context size
    with pytest.raises(pkg.B200FHEError):
        pkg.KeySwitch(ctx, 5, 1, 2)          # n_q not divisible by dnum
    ks = pkg.KeySwitch(ctx, 4, 2, 2)
    assert ks.workspace_bytes() > 0
    with pytest.raises(pkg.B200FHEError):
        ks.apply(0, 0, 0, 0)


@pytest.mark.parametrize("log_n,n_q,n_p,dnum,t", [(10, 6, 2, 3, 65537), (12, 4, 4, 1, 1032193), (13, 8, 4, 2, 65537)])
def test_bgv_keyswitch_parity_and_plaintext_modulus(pkg, O, log_n, n_q, n_p, dnum, t):
    """plain_modulus = t: bit-exact with the oracle's t-scaled composition, and the residual out0 + out1 s - c s' is a
    small multiple of t (the plaintext mod t is untouched by the switch)."""
    import torch
    N, moduli, ctx, tabs = _setup(pkg, O, log_n, n_q, n_p)
    rng = np.random.default_rng(log_n + dnum)
    s, s_new = rng.integers(-1, 2, N), rng.integers(-1, 2, N)
    evk = make_switch_key(O, rng, s, s_new, moduli, n_q, n_p, dnum, tabs, t=t)
    c = O.uniform_limbs(moduli[:n_q], N, seed=21)
    ks = pkg.KeySwitch(ctx, n_q, n_p, dnum, plain_modulus=t)
    eq, ep = _split(evk, n_q)
    d_c, d_eq, d_ep = _dev(torch, c), _dev(torch, eq), _dev(torch, ep)
    d_out = torch.zeros((2, n_q, N), dtype=torch.int64, device="cuda")
    ks.apply(d_c.data_ptr(), d_eq.data_ptr(), d_ep.data_ptr(), d_out.data_ptr())
    torch.cuda.synchronize()
    out = _host(d_out)
    assert np.array_equal(out, O.keyswitch_hybrid(c, evk, moduli, n_q, n_p, dnum, tabs, t=t))
    noise = check_switch_noise(O, out, c, s, s_new, moduli[:n_q], tabs[:n_q], bound_bits=44)
    assert not (noise % t).any()


@pytest.mark.parametrize("log_n,n_q,n_p,dnum", [(10, 6, 2, 3), (12, 8, 4, 2), (13, 4, 4, 1), (17, 48, 16, 3)])
def test_protected_keyswitch_same_bits_and_clean_verdict(pkg, O, log_n, n_q, n_p, dnum):
    """apply_protected (every base conversion checksum-carrying) returns the same bits as apply and an all-zero verdict on
    a fault-free run; the verdict has one entry per checked limb of each of the dnum + 2 conversions."""
    import torch
    N, moduli, ctx, tabs = _setup(pkg, O, log_n, n_q, n_p)
    ks = pkg.KeySwitch(ctx, n_q, n_p, dnum)
    g = torch.Generator(device="cuda").manual_seed(log_n)
    c = torch.randint(0, 1 << 59, (n_q, N), dtype=torch.int64, device="cuda", generator=g)
    eq = torch.randint(0, 1 << 59, (dnum, 2, n_q, N), dtype=torch.int64, device="cuda", generator=g)
    ep = torch.randint(0, 1 << 59, (dnum, 2, n_p, N), dtype=torch.int64, device="cuda", generator=g)
    out = torch.zeros((2, n_q, N), dtype=torch.int64, device="cuda")
    outp = torch.zeros_like(out)
    verdict = torch.full((dnum + 2, n_q + n_p), 7, dtype=torch.int32, device="cuda")
    ks.apply(c.data_ptr(), eq.data_ptr(), ep.data_ptr(), out.data_ptr())
    for _ in range(2):                                          # twice: the two counter sets alternate
        ks.apply_protected(c.data_ptr(), eq.data_ptr(), ep.data_ptr(), outp.data_ptr(), verdict.data_ptr())
        torch.cuda.synchronize()
        assert torch.equal(out, outp)
        assert int(verdict.abs().sum().item()) == 0


@pytest.mark.parametrize("conversion,inject,row_entry", [
    (1, [(2, 5, 100, 1 << 17)], lambda nq, np_, a: a + 5),        # digit 1, contraction output of its target #5
    (0, [(1, 1, 7, 1 << 40)], lambda nq, np_, a: 1),              # digit 0, scale product of source limb 1
    (3, [(3, 0, 9, 1 << 3)], lambda nq, np_, a: 0),               # ModDown of out[0]: y word flipped in memory
    (4, [(2, 2, 0, 1)], lambda nq, np_, a: np_ + 2),              # ModDown of out[1]: lowest bit of one output
])
def test_protected_keyswitch_flags_the_faulty_conversion(pkg, O, conversion, inject, row_entry):
    import torch
    log_n, n_q, n_p, dnum = 11, 6, 2, 3
    N, moduli, ctx, tabs = _setup(pkg, O, log_n, n_q, n_p)
    ks = pkg.KeySwitch(ctx, n_q, n_p, dnum)
    g = torch.Generator(device="cuda").manual_seed(3)
    c = torch.randint(0, 1 << 59, (n_q, N), dtype=torch.int64, device="cuda", generator=g)
    eq = torch.randint(0, 1 << 59, (dnum, 2, n_q, N), dtype=torch.int64, device="cuda", generator=g)
    ep = torch.randint(0, 1 << 59, (dnum, 2, n_p, N), dtype=torch.int64, device="cuda", generator=g)
    clean = torch.zeros((2, n_q, N), dtype=torch.int64, device="cuda")
    out = torch.zeros_like(clean)
    verdict = torch.zeros((dnum + 2, n_q + n_p), dtype=torch.int32, device="cuda")
    ks.apply(c.data_ptr(), eq.data_ptr(), ep.data_ptr(), clean.data_ptr())
    ks.set_inject(conversion, inject)
    ks.apply_protected(c.data_ptr(), eq.data_ptr(), ep.data_ptr(), out.data_ptr(), verdict.data_ptr())
    torch.cuda.synchronize()
    v = verdict.cpu().numpy()
    col = row_entry(n_q, n_p, n_q // dnum)
    assert v[conversion, col] != 0
    v[conversion, col] = 0
    assert not v.any()                                          # only the check of the limb that was hit trips
    assert not torch.equal(out, clean)                          # ... and the fault did corrupt the result
    ks.apply_protected(c.data_ptr(), eq.data_ptr(), ep.data_ptr(), out.data_ptr(), verdict.data_ptr())
    torch.cuda.synchronize()
    assert int(verdict.abs().sum().item()) == 0 and torch.equal(out, clean)      # the injection was consumed


@pytest.mark.parametrize("log_n,n_q,n_p,dnum", [(10, 6, 2, 3), (12, 8, 4, 2), (17, 48, 16, 3)])
def test_fully_protected_keyswitch_also_checks_every_transform(pkg, O, log_n, n_q, n_p, dnum):
    """apply_protected_full: the conversions' checks plus one fused-checksum verdict per limb-transform of the switch (INTT(c), the
    digits' NTTs, ModDown's INTT / NTT): same output bits, all verdicts clean; a fault injected into one conversion trips that
    conversion's check and NONE of the transform checks (the transforms faithfully transform the corrupted data: faults are located)."""
    import torch
    N, moduli, ctx, tabs = _setup(pkg, O, log_n, n_q, n_p)
    ks = pkg.KeySwitch(ctx, n_q, n_p, dnum)
    total, alpha = n_q + n_p, n_q // dnum
    n_ntt = ks.ntt_count()
    assert n_ntt == n_q + dnum * (total - alpha) + 2 * n_p + 2 * n_q
    g = torch.Generator(device="cuda").manual_seed(log_n + 100)
    c = torch.randint(0, 1 << 59, (n_q, N), dtype=torch.int64, device="cuda", generator=g)
    eq = torch.randint(0, 1 << 59, (dnum, 2, n_q, N), dtype=torch.int64, device="cuda", generator=g)
    ep = torch.randint(0, 1 << 59, (dnum, 2, n_p, N), dtype=torch.int64, device="cuda", generator=g)
    out = torch.zeros((2, n_q, N), dtype=torch.int64, device="cuda")
    outp = torch.zeros_like(out)
    ks.apply(c.data_ptr(), eq.data_ptr(), ep.data_ptr(), out.data_ptr())
    for _ in range(2):
        vc = torch.full((dnum + 2, total), 7, dtype=torch.int32, device="cuda")
        vn = torch.full((n_ntt,), 7, dtype=torch.int32, device="cuda")
        ks.apply_protected_full(c.data_ptr(), eq.data_ptr(), ep.data_ptr(), outp.data_ptr(), vc.data_ptr(), vn.data_ptr())
        torch.cuda.synchronize()
        assert torch.equal(out, outp)
        assert int(vc.abs().sum().item()) == 0 and int(vn.abs().sum().item()) == 0
    ks.set_inject(dnum, [(2, 1, 3, 1 << 9)])                    # ModDown of out[0]: one contraction output
    vc.fill_(7); vn.fill_(7)
    ks.apply_protected_full(c.data_ptr(), eq.data_ptr(), ep.data_ptr(), outp.data_ptr(), vc.data_ptr(), vn.data_ptr())
    torch.cuda.synchronize()
    assert int(vc[dnum, n_p + 1].item()) != 0 and int((vc != 0).sum().item()) == 1
    assert int(vn.abs().sum().item()) == 0 and not torch.equal(out, outp)
