"""CPU tier, world_size 2 (and 4) over gloo: the host logic of the limb-sharded ModUp
(fhe_reliability_gpu_b200/modup.py: ownership, gather layout, digit loop) with the device work
replaced -- in this test only -- by an oracle-backed backend.  Every rank's result must equal
the single-process composition INTT -> HPS conversion -> NTT, bit for bit."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class OracleBackend:
    """numpy/torch-CPU stand-in for modup.CudaBackend (test infrastructure)."""

    def __init__(self, O, log_n, moduli):
        self.O, self.log_n, self.N, self.moduli = O, log_n, 1 << log_n, moduli
        self.tabs = [O.Tables(log_n, q) for q in moduli]

    def conv(self, key, src, dst):
        return (list(src), list(dst))

    def empty(self, n_limbs):
        return torch.zeros((n_limbs, self.N), dtype=torch.int64)

    def _u(self, t):
        return t.numpy().view(np.uint64)

    def intt(self, buf, limb_ids):
        for i, g in enumerate(limb_ids):
            self._u(buf)[i] = self.O.ntt_inverse(self._u(buf)[i], self.tabs[g])

    def ntt(self, buf, limb_ids):
        for i, g in enumerate(limb_ids):
            self._u(buf)[i] = self.O.ntt_forward(self._u(buf)[i], self.tabs[g])

    def scale(self, conv, x, y, j0, count):
        src, _ = conv
        p = [self.moduli[g] for g in src]
        for c in range(count):
            j = j0 + c
            h = 1
            for i, pi in enumerate(p):
                if i != j:
                    h = h * pi % p[j]
            inv = pow(h, -1, p[j])
            self._u(y)[c] = np.array([int(v) % p[j] * inv % p[j] for v in self._u(x)[c]], dtype=np.uint64)

    def contract(self, conv, y_full, out, kcount):
        src, dst = conv
        p = [self.moduli[g] for g in src]
        yf = self._u(y_full)
        for k in range(kcount):
            qk = self.moduli[dst[k]]
            acc = [0] * self.N
            for j in range(len(p)):
                hat = 1
                for i, pi in enumerate(p):
                    if i != j:
                        hat = hat * pi % qk
                col = yf[j]
                acc = [(a + int(v) * hat) % qk for a, v in zip(acc, col)]
            self._u(out)[k] = np.array(acc, dtype=np.uint64)

    def all_gather(self, d, send, world):
        out = torch.empty((world * send.shape[0],) + tuple(send.shape[1:]), dtype=send.dtype)
        d.all_gather_into_tensor(out, send)
        return out.reshape((world,) + tuple(send.shape))

    def take_rows(self, flat, rows):
        return flat[rows].contiguous()


def _worker(rank, world, port, log_n, n_q, n_p, dnum, ret):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import oracle as O
        from fhe_reliability_gpu_b200 import modup
        N = 1 << log_n
        moduli = O.primes_create(N, [40] * (n_q + n_p))
        be = OracleBackend(O, log_n, moduli)
        plan = modup.ShardPlan(n_q, n_p, dnum, world, rank)
        a = O.uniform_limbs(moduli[:n_q], N, seed=11)          # same on every rank; each keeps what it owns
        mine = {g: torch.from_numpy(a[g].view(np.int64).copy()) for g in modup.owned(range(n_q), rank, world)}
        ok = True
        for d in range(dnum):
            res = modup.modup_digit(plan, be, dist, d, mine)
            dl, tl = plan.digit_limbs(d), plan.target_limbs(d)
            coeff = np.stack([O.ntt_inverse(a[g], be.tabs[g]) for g in dl])
            conv = O.bconv_hps(coeff, [moduli[g] for g in dl], [moduli[g] for g in tl])
            want = {g: O.ntt_forward(conv[i], be.tabs[g]) for i, g in enumerate(tl)}
            want.update({g: a[g] for g in dl})
            mine_all = modup.owned(range(n_q + n_p), rank, world)
            ok = ok and sorted(res.keys()) == mine_all
            for g in mine_all:
                ok = ok and np.array_equal(res[g].numpy().view(np.uint64), want[g])
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_q,n_p,dnum", [(2, 6, 2, 3), (2, 4, 3, 2), (4, 6, 2, 3), (3, 6, 3, 2)])
def test_sharded_modup_matches_single_process(world, n_q, n_p, dnum):
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = 29500 + (os.getpid() + world * 7 + n_q) % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, 5, n_q, n_p, dnum, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert all(ret.get(r) for r in range(world)), dict(ret)


def test_shard_plan_layout():
    from fhe_reliability_gpu_b200.modup import ShardPlan
    p = ShardPlan(48, 16, 3, 8, 3)                      # C5: L=48, K=16, dnum=3 on 8 GPUs
    assert p.alpha == 16 and p.digit_limbs(1) == list(range(16, 32))
    assert p.my_digit_limbs(1) == [19, 27] and p.gather_counts(1) == ([2] * 8, 2)
    assert p.my_target_limbs(1) == [3, 11, 35, 43, 51, 59]
    assert p.source_order(1)[:4] == [16, 24, 17, 25]
    owners = sorted(g for r in range(8) for g in ShardPlan(48, 16, 3, 8, r).my_target_limbs(1))
    assert owners == p.target_limbs(1)                  # every target limb has exactly one owner
    q = ShardPlan(6, 2, 3, 4, 1)                        # uneven: digit of 2 limbs over 4 ranks
    assert q.gather_counts(0) == ([1, 1, 0, 0], 1) and q.gathered_rows(0) == [0, 1]
    with pytest.raises(ValueError):
        ShardPlan(7, 2, 3, 2, 0)


def test_c_abi_layout_matches_shard_plan():
    """b200fhe_modup_sharded_layout (the C++ plan's ownership / gather order, pure host) == modup.ShardPlan."""
    import fhe_reliability_gpu_b200 as pkg
    from fhe_reliability_gpu_b200 import modup
    for n_q, n_p, dnum in [(48, 16, 3), (12, 4, 3), (6, 2, 2), (4, 2, 2), (9, 3, 3), (8, 1, 1)]:
        for world in (1, 2, 3, 4, 8):
            for rank in range(world):
                own, src, cnt = pkg.binding.modup_sharded_layout(n_q, n_p, dnum, world, rank)
                plan = modup.ShardPlan(n_q, n_p, dnum, world, rank)
                assert own == modup.owned(range(n_q + n_p), rank, world)
                for d in range(dnum):
                    assert src[d] == plan.source_order(d)
                    assert cnt[d] == plan.gather_counts(d)[0]
                    assert sorted(plan.my_digit_limbs(d) + plan.my_target_limbs(d)) == own
