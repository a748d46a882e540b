"""GPU tests of the C ABI's limb-sharded ModUp (b200fhe_shard_* / b200fhe_modup_sharded, config C5).

On ONE device every rank of a world of 1..4 is played in turn through the plan's two local halves (scale / raise) with
the all-gather done by the test (torch.cat in gather order): this covers ownership, the folded scale step, ragged digits,
row placement and the contraction's arbitrary target ranges, bit for bit against the oracle and against the unsharded
b200fhe_keyswitch_modup.  The NCCL call itself runs in the 2-GPU test (skipped on a single-GPU box) and in
`torchrun bench.py --gpus N` (sub-record modup_sharded)."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _oracle_modup(O, c, moduli, n_q, n_p, dnum, tabs):
    total, alpha = n_q + n_p, n_q // dnum
    up = np.zeros((dnum, total, c.shape[1]), dtype=np.uint64)
    for d in range(dnum):
        dl = list(range(d * alpha, (d + 1) * alpha))
        tl = [g for g in range(total) if g not in dl]
        coeff = O.ntt_inverse_batch(c[dl], [tabs[g] for g in dl])
        conv = O.bconv_hps(coeff, [moduli[g] for g in dl], [moduli[g] for g in tl])
        up[d][tl] = O.ntt_forward_batch(conv, [tabs[g] for g in tl])
        up[d][dl] = c[dl]
    return up


@pytest.mark.parametrize("log_n,n_q,n_p,dnum,worlds", [(12, 12, 4, 3, (1, 2, 3, 4)), (13, 6, 2, 2, (1, 2, 4)), (12, 9, 3, 3, (2, 5))])
def test_every_rank_of_an_emulated_world_matches_oracle_and_unsharded(pkg, O, log_n, n_q, n_p, dnum, worlds):
    import torch
    B = pkg.binding
    N, total, alpha = 1 << log_n, n_q + n_p, n_q // dnum
    moduli = pkg.primes_create(N, [60] * total)
    rng = np.random.default_rng(log_n * 100 + n_q)
    c = np.stack([rng.integers(0, q, size=N, dtype=np.uint64) for q in moduli[:n_q]])
    tabs = [O.Tables(log_n, q) for q in moduli]
    want = _oracle_modup(O, c, moduli, n_q, n_p, dnum, tabs)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).cuda()
    ctx = pkg.Context(log_n, moduli)
    ks = pkg.KeySwitch(ctx, n_q, n_p, dnum)
    d_ref = torch.zeros((dnum, total, N), dtype=torch.int64, device="cuda")
    ks.modup(dev(c).data_ptr(), d_ref.data_ptr())
    torch.cuda.synchronize()
    assert np.array_equal(d_ref.cpu().numpy().view(np.uint64), want)
    for world in worlds:
        shards = [B.Shard(r, world, 0, None) for r in range(world)]
        plans = [B.ModUpSharded(s, log_n, moduli, n_q, n_p, dnum) for s in shards]
        d_c, d_y = [], []
        for r, p in enumerate(plans):
            own_q = [g for g in p.own_limbs if g < n_q]
            assert p.own_limbs == [g for g in range(total) if g % world == r] and p.n_own_q == len(own_q)
            d_c.append(dev(c[own_q]) if own_q else torch.zeros((1, N), dtype=torch.int64, device="cuda"))
            d_y.append(torch.zeros_like(d_c[-1]))
            p.scale(d_c[-1].data_ptr(), d_y[-1].data_ptr())
        # the "all-gather": digit d = every rank's scaled rows of that digit, rank-major
        y_full = []
        for d in range(dnum):
            parts = []
            for r, p in enumerate(plans):
                own_q = [g for g in p.own_limbs if g < n_q]
                rows = [i for i, g in enumerate(own_q) if d * alpha <= g < (d + 1) * alpha]
                if rows:
                    parts.append(d_y[r][rows[0]:rows[-1] + 1])
            y_full.append(torch.cat(parts).contiguous())
            assert y_full[-1].shape[0] == alpha
        for r, p in enumerate(plans):
            d_up = torch.zeros((dnum, p.n_own_total, N), dtype=torch.int64, device="cuda")
            p.raise_(d_c[r].data_ptr(), [t.data_ptr() for t in y_full], d_up.data_ptr())
            torch.cuda.synchronize()
            assert np.array_equal(d_up.cpu().numpy().view(np.uint64), want[:, p.own_limbs]), (world, r)
            if world > 1:
                with pytest.raises(B.B200FHEError):                       # no communicator on an own-transport shard
                    p.run(d_c[r].data_ptr(), d_up.data_ptr())
        for p in plans:
            p.close()
        for s in shards:
            s.close()
    ks.close()


def test_world_of_one_is_the_unsharded_modup(pkg, O):
    import torch
    B = pkg.binding
    log_n, n_q, n_p, dnum = 14, 12, 4, 3
    N, total = 1 << log_n, n_q + n_p
    moduli = pkg.primes_create(N, [60] * total)
    rng = np.random.default_rng(77)
    c = np.stack([rng.integers(0, q, size=N, dtype=np.uint64) for q in moduli[:n_q]])
    d_c = torch.from_numpy(c.view(np.int64)).cuda()
    shard = B.Shard(0, 1, 0, None)
    plan = B.ModUpSharded(shard, log_n, moduli, n_q, n_p, dnum)
    d_up = torch.zeros((dnum, total, N), dtype=torch.int64, device="cuda")
    plan.run(d_c.data_ptr(), d_up.data_ptr())
    ctx = pkg.Context(log_n, moduli)
    ks = pkg.KeySwitch(ctx, n_q, n_p, dnum)
    d_ref = torch.zeros_like(d_up)
    ks.modup(d_c.data_ptr(), d_ref.data_ptr())
    torch.cuda.synchronize()
    assert torch.equal(d_up, d_ref)
    tabs = [O.Tables(log_n, q) for q in moduli]
    assert np.array_equal(d_up.cpu().numpy().view(np.uint64), _oracle_modup(O, c, moduli, n_q, n_p, dnum, tabs))
    plan.close(); shard.close(); ks.close()


def test_two_gpus_over_nccl_bit_identical_to_unsharded():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (the single-GPU emulation above covers everything but the NCCL call)")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29534", os.path.join(root, "bench_modup.py"),
                        "--log-n", "14", "--limbs", "12", "--special", "4", "--steps", "2", "--no-graph"],
                       capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    assert "bit-exact" in p.stdout
