"""b200fhe_bconv_modup_static (the allocation-free ModUp base conversion: scratch in a file-scope static __device__ array,
no counterpart in the reference) against the existing b200fhe_bconv_modup, through the C ABI, on two calls in a row with
different inputs (the array keeps the previous call's contents), bit-exact as every integer test here.  Collected last on
purpose (file name): it must pass alone in a fresh process and after the other tests."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_static_scratch_modup_matches_existing_entry_point_twice(pkg, O):
    import torch
    log_n, m, K = 14, 16, 48
    N = 1 << log_n
    moduli = pkg.primes_create(N, [60] * (m + K))
    ctx = pkg.Context(log_n, moduli)
    bc = pkg.BaseConv(ctx, list(range(m)), list(range(m, m + K)))
    d_ref = torch.zeros((K, N), dtype=torch.int64, device="cuda")
    d_out = torch.zeros((K, N), dtype=torch.int64, device="cuda")
    for call in range(2):
        x = O.uniform_limbs(moduli[:m], N, seed=900 + call)
        d_in = torch.from_numpy(x.view(np.int64)).cuda()
        d_ref.zero_(); d_out.zero_()
        bc.modup(d_in.data_ptr(), d_ref.data_ptr())
        bc.modup_static(d_in.data_ptr(), d_out.data_ptr())
        torch.cuda.synchronize()
        assert torch.equal(d_out, d_ref), "call %d: static-scratch ModUp differs from b200fhe_bconv_modup" % call
        if call == 0:
            assert np.array_equal(d_out.cpu().numpy().view(np.uint64), O.bconv_hps(x, moduli[:m], moduli[m:]))
    # a smaller plan afterwards reuses the front of the same array
    bc2 = pkg.BaseConv(ctx, [3, 5, 7], [0, 1, 2, 40])
    x = O.uniform_limbs([moduli[i] for i in (3, 5, 7)], N, seed=77)
    d_in = torch.from_numpy(x.view(np.int64)).cuda()
    a = torch.zeros((4, N), dtype=torch.int64, device="cuda"); b = torch.zeros_like(a)
    bc2.modup(d_in.data_ptr(), a.data_ptr()); bc2.modup_static(d_in.data_ptr(), b.data_ptr())
    torch.cuda.synchronize()
    assert torch.equal(a, b)


def test_static_scratch_limit_is_enforced_with_a_clear_error(pkg):
    import torch
    log_n, m = 16, 16                                    # 16 * 65536 words = 8 MiB > the 4 MiB array
    N = 1 << log_n
    moduli = pkg.primes_create(N, [60] * (m + 2))
    ctx = pkg.Context(log_n, moduli)
    bc = pkg.BaseConv(ctx, list(range(m)), [m, m + 1])
    d_in = torch.zeros((m, N), dtype=torch.int64, device="cuda"); d_out = torch.zeros((2, N), dtype=torch.int64, device="cuda")
    with pytest.raises(pkg.B200FHEError, match="exceeds the static scratch"):
        bc.modup_static(d_in.data_ptr(), d_out.data_ptr())
    bc8 = pkg.BaseConv(ctx, list(range(8)), [m, m + 1])    # 8 * 65536 words = exactly the limit: accepted
    bc8.modup_static(d_in.data_ptr(), d_out.data_ptr())
    torch.cuda.synchronize()
    exact = pkg.BaseConv(ctx, [0, 1], [2], mode=1)
    with pytest.raises(pkg.B200FHEError):
        exact.modup_static(d_in.data_ptr(), d_out.data_ptr())
