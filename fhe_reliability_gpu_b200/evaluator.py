"""Device-side evaluator steps with the names reliability_test/dotprod_test.cu uses (arithmetic only):

  dotprod_test.cu:113  multiply(context, a, b)              -> Evaluator.multiply
  dotprod_test.cu:114  relinearize_inplace(ct, relin_keys)  -> Evaluator.relinearize
  dotprod_test.cu:115  mod_switch_to_next_inplace(ct)       -> Evaluator.rescale_to_next   (division by the last prime)
  dotprod_test.cu:146  rotate_inplace(ct, step, galois_keys)-> Evaluator.rotate_galois
  dotprod_test.cu:150  add_inplace(a, b)                    -> Evaluator.add_inplace

Ciphertexts are torch int64 CUDA tensors [size][L][N] in NTT form over the context's first n_q moduli (the Phantom
device layout, dotprod_test.cu:42-46); keys are (evk_q [dnum][2][n_q][N], evk_p [dnum][2][n_p][N]) tensors.  Encoding,
encryption, key generation and the BGV plaintext-modulus corrections stay on the caller's side (out of scope, DESIGN.md
section 9).  Every step is CUDA work behind include/b200fhe.h; there is no CPU fallback."""
from . import binding as B


class Evaluator:
    def __init__(self, ctx, n_q, n_p, dnum, plain_modulus=1):
        """ctx: Context over n_q ciphertext moduli followed by n_p special primes.  plain_modulus t > 1 selects the BGV
        flavour of ModDown / rescale (the subtracted term is made a multiple of t, so plaintexts mod t survive)."""
        import torch
        self.torch = torch
        self.ctx, self.n_q, self.n_p, self.dnum, self.N = ctx, n_q, n_p, dnum, ctx.N
        self.t = plain_modulus
        self.ks = B.KeySwitch(ctx, n_q, n_p, dnum, plain_modulus)
        self._rescale = None

    def _stream(self):
        return self.torch.cuda.current_stream().cuda_stream

    def _check(self, ct, size=None):
        if ct.dtype != self.torch.int64 or not ct.is_cuda or not ct.is_contiguous():
            raise ValueError("ciphertext must be a contiguous int64 CUDA tensor")
        if ct.dim() != 3 or ct.shape[1] != self.n_q or ct.shape[2] != self.N or (size and ct.shape[0] != size):
            raise ValueError("ciphertext shape %s, expected [%s][%d][%d]" % (tuple(ct.shape), size or "size", self.n_q, self.N))

    def multiply(self, a, b):
        """(a0, a1) x (b0, b1) -> (a0 b0, a0 b1 + a1 b0, a1 b1)."""
        self._check(a, 2); self._check(b, 2)
        out = self.torch.empty((3, self.n_q, self.N), dtype=self.torch.int64, device=a.device)
        B.check(B.lib().b200fhe_ct_tensor(self.ctx._h, a.data_ptr(), b.data_ptr(), out.data_ptr(), self.n_q, 0, self._stream()))
        return out

    def add_inplace(self, a, b):
        self._check(a); self._check(b, a.shape[0])
        B.check(B.lib().b200fhe_poly_add(self.ctx._h, a.data_ptr(), b.data_ptr(), a.shape[0], self.n_q, 0, self._stream()))
        return a

    def sub_inplace(self, a, b):
        self._check(a); self._check(b, a.shape[0])
        B.check(B.lib().b200fhe_poly_sub(self.ctx._h, a.data_ptr(), b.data_ptr(), a.shape[0], self.n_q, 0, self._stream()))
        return a

    def switch_key(self, poly, key):
        """poly [n_q][N] under the key's source secret -> [2][n_q][N] under the target secret."""
        evk_q, evk_p = key
        out = self.torch.empty((2, self.n_q, self.N), dtype=self.torch.int64, device=poly.device)
        self.ks.apply(poly.data_ptr(), evk_q.data_ptr(), evk_p.data_ptr(), out.data_ptr(), stream=self._stream())
        return out

    def relinearize(self, ct3, relin_key):
        """(d0, d1, d2) -> (d0, d1) + KeySwitch_{s^2 -> s}(d2)."""
        self._check(ct3, 3)
        out = self.switch_key(ct3[2], relin_key)
        return self.add_inplace(out, ct3[:2].contiguous())

    def apply_galois(self, ct, galois_elt):
        """x -> x^galois_elt on every polynomial of ct (still under s(x^galois_elt))."""
        self._check(ct)
        out = self.torch.empty_like(ct)
        B.check(B.lib().b200fhe_galois_ntt(self.ctx._h, ct.data_ptr(), out.data_ptr(), galois_elt, ct.shape[0] * self.n_q, self._stream()))
        return out

    def rotate_galois(self, ct, galois_elt, galois_key):
        """rotate_inplace: automorphism, then switch c1 back from s(x^g) to s."""
        self._check(ct, 2)
        g = self.apply_galois(ct, galois_elt)
        out = self.switch_key(g[1], galois_key)
        B.check(B.lib().b200fhe_poly_add(self.ctx._h, out.data_ptr(), g.data_ptr(), 1, self.n_q, 0, self._stream()))
        return out

    def rescale_to_next(self, ct, q_ctx):
        """Drop the last ciphertext prime: [size][n_q][N] -> [size][n_q - 1][N] = floor(ct / q_last) (no rounding term);
        with a plain modulus t: (ct - t [ct t^-1]_q_last) / q_last, BGV's mod_switch_to_next (the plaintext becomes
        m q_last^-1 mod t, the correction factor BGV libraries track).
        q_ctx: a Context over exactly the n_q ciphertext moduli (the last one plays the special prime of a ModDown)."""
        self._check(ct)
        if self._rescale is None or self._rescale[0] is not q_ctx:
            self._rescale = (q_ctx, B.KeySwitch(q_ctx, self.n_q - 1, 1, 0, self.t))
        plan = self._rescale[1]
        size, nq1 = ct.shape[0], self.n_q - 1
        out = self.torch.empty((size, nq1, self.N), dtype=self.torch.int64, device=ct.device)
        for p0 in range(0, size, 2):
            n = min(2, size - p0)
            acc_q = ct[p0:p0 + n, :nq1].contiguous()
            acc_p = ct[p0:p0 + n, nq1:].contiguous()
            plan.moddown(acc_q.data_ptr(), acc_p.data_ptr(), out[p0].data_ptr(), n, stream=self._stream())
        return out
