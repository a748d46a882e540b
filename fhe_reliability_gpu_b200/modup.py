"""Limb-sharded key-switch ModUp (config C5): INTT -> base conversion -> NTT with the
digit's limbs all-gathered over NCCL/NVLink between the two conversion steps.

The reference has no multi-GPU code; this is the B200-native shape of the
step that Phantom's relinearize_inplace / rotate_inplace run on one device
(reliability_test/dotprod_test.cu:114,146; share of key-switch time at dnum=3:
profile_framewk/draw_dnum.py:5-9).

Layout: the ciphertext polynomial lives in the NTT domain as [L_total][N] limbs over the
moduli Q (L of them) followed by the special primes P (K of them).  Limb g (global index)
is owned by rank g % world.  Digit d = limbs [d*alpha, (d+1)*alpha) of Q.  For one digit:

  1. every rank runs the inverse NTT on the digit limbs it owns           (local)
  2. ... and scales them: y_j = x_j (D/p_j)^-1 mod p_j                     (local, b200fhe_bconv_scale)
  3. all-gather of the alpha scaled limbs                                  (the ONE collective)
  4. every rank contracts the gathered digit into the target limbs it owns (local, b200fhe_bconv_contract)
  5. ... and runs the forward NTT on them                                  (local)

Digit limbs themselves are copied through unchanged (they already are the residues mod
their own moduli).  `Backend` isolates the device work so that the host logic (ownership
maps, gather layout, digit loop) is testable on CPU with world_size 2 over gloo
(tests/test_modup_sharded_cpu.py drives it with an oracle-backed backend); the product
backend below is CUDA-only and raises without a device.
"""
import numpy as np


def owner(limb, world):
    return limb % world


def owned(limbs, rank, world):
    return [g for g in limbs if g % world == rank]


class ShardPlan:
    """Ownership and gather layout of one digit: pure host arithmetic."""

    def __init__(self, n_q, n_p, dnum, world, rank):
        if n_q % dnum:
            raise ValueError("L=%d not divisible by dnum=%d" % (n_q, dnum))
        self.n_q, self.n_p, self.dnum, self.world, self.rank = n_q, n_p, dnum, world, rank
        self.alpha = n_q // dnum
        self.total = n_q + n_p

    def digit_limbs(self, d):
        return list(range(d * self.alpha, (d + 1) * self.alpha))

    def target_limbs(self, d):
        dl = set(self.digit_limbs(d))
        return [g for g in range(self.total) if g not in dl]

    def my_digit_limbs(self, d):
        return owned(self.digit_limbs(d), self.rank, self.world)

    def my_target_limbs(self, d):
        return owned(self.target_limbs(d), self.rank, self.world)

    def gather_counts(self, d):
        """limbs each rank contributes to the all-gather of digit d (equal counts are required by
        all_gather_into_tensor; uneven digits are padded to the maximum)."""
        c = [len(owned(self.digit_limbs(d), r, self.world)) for r in range(self.world)]
        return c, max(c)

    def source_order(self, d):
        """digit limbs in gather order (rank-major): the order of the conversion's source base, so the
        all-gathered buffer IS the scaled digit and every rank's contribution is one contiguous run."""
        return [g for r in range(self.world) for g in owned(self.digit_limbs(d), r, self.world)]

    def gathered_rows(self, d):
        """rows of the gathered [world][cmax] buffer that hold real limbs (padding rows skipped)."""
        counts, cmax = self.gather_counts(d)
        return [r * cmax + i for r in range(self.world) for i in range(counts[r])]


class CudaBackend:
    """Device work of the sharded ModUp through the C ABI (torch only for memory, streams, NCCL).

    The rank's context lists the limbs it owns FIRST (ascending), so a set of owned limbs is a few
    contiguous runs of modulus indices and each run is one batched kernel launch."""

    def __init__(self, pkg, log_n, moduli, rank, world, device):
        import torch
        self.torch = torch
        self.pkg = pkg
        self.N = 1 << log_n
        total = len(moduli)
        self.order = owned(range(total), rank, world) + [g for g in range(total) if g % world != rank]
        self.pos = {g: i for i, g in enumerate(self.order)}
        self.ctx = pkg.Context(log_n, [moduli[g] for g in self.order], device=device)
        self.device = torch.device("cuda", device)
        self.convs = {}

    def conv(self, key, src, dst):
        if key not in self.convs:
            self.convs[key] = self.pkg.BaseConv(self.ctx, [self.pos[g] for g in src], [self.pos[g] for g in dst])
        return self.convs[key]

    def empty(self, n_limbs):
        # every row that is read later is written first (padding rows of a gather are skipped by gathered_rows)
        return self.torch.empty((n_limbs, self.N), dtype=self.torch.int64, device=self.device)

    def stack(self, limbs, rows):
        """[rows][N] buffer whose first len(limbs) rows are the given limbs: one copy kernel instead of one per limb."""
        x = self.empty(rows)
        if limbs:
            self.torch.stack(limbs, out=x[:len(limbs)])
        return x

    def _runs(self, limb_ids):
        runs, i = [], 0
        while i < len(limb_ids):
            j = i
            while j + 1 < len(limb_ids) and self.pos[limb_ids[j + 1]] == self.pos[limb_ids[j]] + 1:
                j += 1
            runs.append((i, j - i + 1, self.pos[limb_ids[i]]))
            i = j + 1
        return runs

    def intt(self, buf, limb_ids):
        s = self.torch.cuda.current_stream().cuda_stream
        for i0, cnt, start in self._runs(limb_ids):
            self.ctx.nwt_backward(buf[i0].data_ptr(), 1, L=cnt, start_mod=start, stream=s)

    def ntt(self, buf, limb_ids):
        s = self.torch.cuda.current_stream().cuda_stream
        for i0, cnt, start in self._runs(limb_ids):
            self.ctx.nwt_forward(buf[i0].data_ptr(), 1, L=cnt, start_mod=start, stream=s)

    def scale(self, conv, x, y, j0, count):
        if count:
            conv.scale(x.data_ptr(), y.data_ptr(), j0, count, stream=self.torch.cuda.current_stream().cuda_stream)

    def contract(self, conv, y_full, out, kcount):
        if kcount:
            conv.contract(y_full.data_ptr(), out.data_ptr(), 0, kcount, stream=self.torch.cuda.current_stream().cuda_stream)

    def all_gather(self, dist, send, world):
        if world == 1:
            return send.unsqueeze(0)
        out = self.torch.empty((world * send.shape[0],) + tuple(send.shape[1:]), dtype=send.dtype, device=send.device)
        dist.all_gather_into_tensor(out, send)      # concatenation along dim 0 = [world][cmax][N]
        return out.reshape((world,) + tuple(send.shape))

    def take_rows(self, flat, rows):
        return flat.index_select(0, self.torch.tensor(rows, dtype=self.torch.long, device=flat.device)).contiguous()


def modup_digit(plan, backend, dist, d, my_limbs):
    """One digit of the sharded ModUp.

    my_limbs: {global limb index: [N] tensor} with the NTT-domain limbs of Q this rank owns (at least
    those of digit d).  Returns {global limb index: [N]} for every limb of Q u P this rank owns: the
    digit raised to the full basis, NTT domain (the digit's own limbs pass through)."""
    world = plan.world
    src = plan.source_order(d)
    mine_d, mine_t = plan.my_digit_limbs(d), plan.my_target_limbs(d)
    counts, cmax = plan.gather_counts(d)
    conv = backend.conv((d, plan.rank), src, mine_t) if mine_t else None
    sconv = conv if conv is not None else backend.conv((d, plan.rank, "s"), src, src[:1])
    # 1-2: inverse NTT and scaling of the owned digit limbs (one contiguous run of the source order)
    if hasattr(backend, "stack"):
        x = backend.stack([my_limbs[g] for g in mine_d], max(cmax, 1))
    else:
        x = backend.empty(max(cmax, 1))
        for i, g in enumerate(mine_d):
            x[i] = my_limbs[g]
    backend.intt(x[:len(mine_d)], mine_d)
    y = backend.empty(max(cmax, 1))
    backend.scale(sconv, x, y, src.index(mine_d[0]) if mine_d else 0, len(mine_d))
    # 3: the one collective of the path
    gathered = backend.all_gather(dist, y, world)                       # [world][cmax][N]
    flat = gathered.reshape(world * max(cmax, 1), -1)
    y_full = flat if min(counts) == cmax else backend.take_rows(flat, plan.gathered_rows(d))
    # 4-5: contraction into, and forward NTT of, the owned target limbs
    out = backend.empty(max(len(mine_t), 1))
    backend.contract(conv, y_full, out, len(mine_t))
    backend.ntt(out[:len(mine_t)], mine_t)
    res = {g: out[i] for i, g in enumerate(mine_t)}
    for g in mine_d:
        res[g] = my_limbs[g]
    return res
