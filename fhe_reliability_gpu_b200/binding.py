"""ctypes binding of lib/libb200fhe.so (include/b200fhe.h).

This is the host-side mirror the Python prototypes/scripts call.  There is no
CPU fallback: if the shared library is missing, or no CUDA device is usable,
every compute entry raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200FHE_LIB") or os.path.join(_HERE, "lib", "libb200fhe.so")   # env override: A/B builds
_lib = None

u64, i32, sz, vp = C.c_uint64, C.c_int, C.c_size_t, C.c_void_p
U64P = C.POINTER(C.c_uint64)

# name -> (restype, argtypes); MUST list every symbol include/b200fhe.h declares
SIGNATURES = {
    "b200fhe_version": (i32, []),
    "b200fhe_last_error": (C.c_char_p, []),
    "b200fhe_primes_create": (i32, [u64, C.POINTER(i32), i32, U64P]),
    "b200fhe_min_primitive_root": (i32, [u64, u64, U64P]),
    "b200fhe_host_tables": (i32, [i32, u64, U64P, U64P, U64P, U64P, U64P, U64P]),
    "b200fhe_host_tables_psi": (i32, [i32, u64, u64, U64P, U64P, U64P, U64P, U64P, U64P]),
    "b200fhe_const_ratio": (i32, [u64, U64P]),
    "b200fhe_ctx_create": (i32, [i32, U64P, i32, i32, C.POINTER(vp)]),
    "b200fhe_ctx_create_empty": (i32, [i32, i32, i32, C.POINTER(vp)]),
    "b200fhe_ctx_set_tables": (i32, [vp, i32, u64, U64P, U64P, U64P, U64P, u64, u64, vp]),
    "b200fhe_ctx_destroy": (i32, [vp]),
    "b200fhe_ctx_info": (i32, [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]),
    "b200fhe_ctx_moduli": (i32, [vp, U64P]),
    "b200fhe_nwt_forward": (i32, [vp, vp, sz, sz, sz, vp]),
    "b200fhe_nwt_backward": (i32, [vp, vp, sz, sz, sz, vp]),
    "b200fhe_nwt_forward_host": (i32, [vp, U64P, sz, sz, sz]),
    "b200fhe_nwt_backward_host": (i32, [vp, U64P, sz, sz, sz]),
    "b200fhe_nwt_forward_protected": (i32, [vp, vp, sz, sz, sz, vp, vp, vp, sz, vp]),
    "b200fhe_nwt_forward_protected_host": (i32, [vp, U64P, sz, sz, sz, C.POINTER(C.c_uint32), U64P, vp, sz]),
    "b200fhe_nwt_backward_protected": (i32, [vp, vp, sz, sz, sz, vp, vp, vp]),
    "b200fhe_nwt_backward_protected_host": (i32, [vp, U64P, sz, sz, sz, C.POINTER(C.c_uint32), U64P]),
    "b200fhe_abft_weights": (i32, [vp, i32, U64P, U64P]),
    "b200fhe_four_step_protected_host": (i32, [u64, u64, sz, U64P, U64P, sz, U64P, C.POINTER(C.c_uint32), vp, sz, i32]),
    "b200fhe_relia_op_ranges": (i32, [sz, U64P, U64P, U64P]),
    "b200fhe_relia_four_step_host": (i32, [u64, u64, sz, i32, U64P, U64P, U64P, sz, vp, sz, C.POINTER(C.c_uint32), i32]),
    "b200fhe_relia_standard_host": (i32, [u64, u64, sz, U64P, U64P, U64P, sz, vp, sz, C.POINTER(C.c_uint32), i32]),
    "b200fhe_pmult_accum": (i32, [vp, C.POINTER(vp), C.POINTER(vp), sz, vp, sz, sz, sz, i32, vp]),
    "b200fhe_pmult_accum_host": (i32, [vp, U64P, U64P, sz, U64P, sz, sz, sz]),
    "b200fhe_bconv_create": (i32, [vp, C.POINTER(i32), i32, C.POINTER(i32), i32, C.POINTER(vp)]),
    "b200fhe_bconv_create_scaled": (i32, [vp, C.POINTER(i32), i32, C.POINTER(i32), i32, u64, C.POINTER(vp)]),
    "b200fhe_bconv_create_mode": (i32, [vp, C.POINTER(i32), i32, C.POINTER(i32), i32, i32, C.POINTER(vp)]),
    "b200fhe_bconv_destroy": (i32, [vp]),
    "b200fhe_bconv_scale": (i32, [vp, vp, vp, sz, sz, vp]),
    "b200fhe_bconv_contract": (i32, [vp, vp, vp, sz, sz, vp]),
    "b200fhe_bconv_modup": (i32, [vp, vp, vp, vp]),
    "b200fhe_bconv_modup_static": (i32, [vp, vp, vp, vp]),
    "b200fhe_bconv_modup_host": (i32, [vp, U64P, U64P]),
    "b200fhe_bconv_modup_protected": (i32, [vp, vp, vp, vp, vp, vp, sz, vp]),
    "b200fhe_bconv_modup_protected_host": (i32, [vp, U64P, U64P, C.POINTER(C.c_uint32), U64P, vp, sz]),
    "b200fhe_keyswitch_create": (i32, [vp, i32, i32, i32, C.POINTER(vp)]),
    "b200fhe_keyswitch_create_bgv": (i32, [vp, i32, i32, i32, u64, C.POINTER(vp)]),
    "b200fhe_keyswitch_destroy": (i32, [vp]),
    "b200fhe_keyswitch_info": (i32, [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32), C.POINTER(sz)]),
    "b200fhe_keyswitch_modup": (i32, [vp, vp, vp, vp]),
    "b200fhe_keyswitch_moddown": (i32, [vp, vp, vp, vp, sz, vp]),
    "b200fhe_keyswitch_apply": (i32, [vp, vp, vp, vp, vp, vp]),
    "b200fhe_keyswitch_apply_protected": (i32, [vp, vp, vp, vp, vp, vp, vp]),
    "b200fhe_keyswitch_apply_protected_full": (i32, [vp, vp, vp, vp, vp, vp, vp, vp]),
    "b200fhe_keyswitch_ntt_count": (i32, [vp, C.POINTER(sz)]),
    "b200fhe_keyswitch_set_inject": (i32, [vp, i32, vp, sz]),
    "b200fhe_shard_unique_id": (i32, [C.POINTER(C.c_uint8)]),
    "b200fhe_shard_init": (i32, [vp, i32, i32, C.POINTER(vp)]),
    "b200fhe_shard_init_id": (i32, [C.POINTER(C.c_uint8), i32, i32, i32, C.POINTER(vp)]),
    "b200fhe_shard_destroy": (i32, [vp]),
    "b200fhe_shard_info": (i32, [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]),
    "b200fhe_modup_sharded_layout": (i32, [i32, i32, i32, i32, i32, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]),
    "b200fhe_modup_sharded_create": (i32, [vp, i32, U64P, i32, i32, i32, C.POINTER(vp)]),
    "b200fhe_modup_sharded_destroy": (i32, [vp]),
    "b200fhe_modup_sharded_info": (i32, [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]),
    "b200fhe_modup_sharded": (i32, [vp, vp, vp, vp]),
    "b200fhe_modup_sharded_scale": (i32, [vp, vp, vp, vp]),
    "b200fhe_modup_sharded_raise": (i32, [vp, vp, C.POINTER(vp), vp, vp]),
    "b200fhe_bgv_create": (i32, [i32, U64P, i32, i32, u64, u64, i32, C.POINTER(vp)]),
    "b200fhe_bgv_destroy": (i32, [vp]),
    "b200fhe_bgv_info": (i32, [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32), U64P, U64P]),
    "b200fhe_bgv_gen_secretkey": (i32, [vp]),
    "b200fhe_bgv_gen_publickey": (i32, [vp]),
    "b200fhe_bgv_encode": (i32, [vp, U64P, sz, U64P]),
    "b200fhe_bgv_decode": (i32, [vp, U64P, U64P]),
    "b200fhe_bgv_encrypt": (i32, [vp, U64P, vp]),
    "b200fhe_bgv_decrypt": (i32, [vp, vp, i32, i32, u64, U64P]),
    "b200fhe_bgv_multiply": (i32, [vp, vp, vp, vp, i32]),
    "b200fhe_bgv_relinearize": (i32, [vp, vp, i32]),
    "b200fhe_bgv_mod_switch_to_next": (i32, [vp, vp, i32, i32, vp, U64P]),
    "b200fhe_bgv_rotate": (i32, [vp, vp, i32, i32, i32]),
    "b200fhe_bgv_add": (i32, [vp, vp, vp, i32, i32]),
    "b200fhe_ct_tensor": (i32, [vp, vp, vp, vp, sz, sz, vp]),
    "b200fhe_poly_add": (i32, [vp, vp, vp, sz, sz, sz, vp]),
    "b200fhe_poly_sub": (i32, [vp, vp, vp, sz, sz, sz, vp]),
    "b200fhe_galois_ntt": (i32, [vp, vp, vp, C.c_uint32, sz, vp]),
    "b200fhe_barrett_params": (i32, [u64, C.POINTER(i32), U64P]),
    "b200fhe_barrett_fold": (i32, [u64, vp, vp, vp, sz, sz, i32, i32, i32, i32, vp, sz, vp, vp, vp, vp]),
    "b200fhe_barrett_fold_host": (i32, [u64, U64P, U64P, U64P, sz, sz, i32, i32, i32, i32, vp, sz, U64P, U64P, C.POINTER(C.c_uint32), i32]),
    "b200fhe_flip_bit": (i32, [vp, sz, C.c_uint, vp]),
    "b200fhe_host_alloc": (i32, [C.POINTER(vp), sz]),
    "b200fhe_host_free": (i32, [vp]),
    "b200fhe_debug_nwt_pass": (i32, [vp, vp, sz, sz, sz, i32, i32, vp]),
    "b200fhe_launch_count": (C.c_ulonglong, []),
}


class B200FHEError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200FHEError(
                "%s not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(no CPU fallback exists)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise B200FHEError("b200fhe error %d: %s" % (rc, lib().b200fhe_last_error().decode()))


def _p(a):
    assert a.dtype == np.uint64 and a.flags.c_contiguous
    return a.ctypes.data_as(U64P)


def primes_create(N, bits):
    bits = list(bits)
    out = np.zeros(len(bits), dtype=np.uint64)
    check(lib().b200fhe_primes_create(N, (i32 * len(bits))(*bits), len(bits), _p(out)))
    return [int(x) for x in out]


def min_primitive_root(N, q):
    r = u64(0)
    check(lib().b200fhe_min_primitive_root(N, q, C.byref(r)))
    return int(r.value)


def host_tables(log_n, q):
    N = 1 << log_n
    arrs = [np.zeros(N, dtype=np.uint64) for _ in range(4)]
    ni, nis = u64(0), u64(0)
    check(lib().b200fhe_host_tables(log_n, q, *[_p(a) for a in arrs], C.byref(ni), C.byref(nis)))
    return arrs + [int(ni.value), int(nis.value)]


class Context:
    """Moduli + tables resident on one GPU (b200fhe_ctx)."""

    def __init__(self, log_n, moduli, device=0):
        self.log_n, self.N = log_n, 1 << log_n
        self.moduli = [int(q) for q in moduli]
        self.L = len(self.moduli)
        self.device = device
        arr = np.array(self.moduli, dtype=np.uint64)
        h = vp()
        check(lib().b200fhe_ctx_create(log_n, _p(arr), self.L, device, C.byref(h)))
        self._h = h

    @classmethod
    def from_psi(cls, log_n, moduli, psis, device=0):
        """Context whose limb i uses the caller's primitive 2N-th root psis[i] instead of the minimal
        one (DNTTTable::init + set with caller-built tables, reliability_test/ntt_test.cu:57-68)."""
        self = cls.__new__(cls)
        self.log_n, self.N = log_n, 1 << log_n
        self.moduli = [int(q) for q in moduli]
        self.L = len(self.moduli)
        self.device = device
        h = vp()
        check(lib().b200fhe_ctx_create_empty(log_n, self.L, device, C.byref(h)))
        self._h = h
        for i, (q, psi) in enumerate(zip(self.moduli, psis)):
            arrs = [np.zeros(self.N, dtype=np.uint64) for _ in range(4)]
            ni, nis = u64(0), u64(0)
            check(lib().b200fhe_host_tables_psi(log_n, q, int(psi), *[_p(a) for a in arrs], C.byref(ni), C.byref(nis)))
            check(lib().b200fhe_ctx_set_tables(h, i, q, *[_p(a) for a in arrs], ni.value, nis.value, None))
        return self

    def close(self):
        if getattr(self, "_h", None):
            lib().b200fhe_ctx_destroy(self._h)
            self._h = None

    __del__ = close

    # ---- device-pointer entry points (ptr: int device address, stream: int cudaStream_t)
    def nwt_forward(self, ptr, n_poly, L=None, start_mod=0, stream=0):
        check(lib().b200fhe_nwt_forward(self._h, ptr, n_poly, self.L if L is None else L, start_mod, stream))

    def nwt_backward(self, ptr, n_poly, L=None, start_mod=0, stream=0):
        check(lib().b200fhe_nwt_backward(self._h, ptr, n_poly, self.L if L is None else L, start_mod, stream))

    # ---- host-buffer entry points (numpy uint64 [n_poly][L][N], in place)
    def _host(self, fn, a, L, start_mod):
        L = self.L if L is None else L
        assert a.dtype == np.uint64 and a.flags.c_contiguous and a.size % (L * self.N) == 0
        check(fn(self._h, _p(a), a.size // (L * self.N), L, start_mod))
        return a

    def nwt_forward_host(self, a, L=None, start_mod=0):
        return self._host(lib().b200fhe_nwt_forward_host, a, L, start_mod)

    def nwt_backward_host(self, a, L=None, start_mod=0):
        return self._host(lib().b200fhe_nwt_backward_host, a, L, start_mod)


class Inject(C.Structure):
    """b200fhe_inject: one fault in relia_ntt_sim.cpp's operation numbering."""
    _fields_ = [("limb", C.c_uint32), ("stage", C.c_uint32), ("butterfly", C.c_uint64), ("kind", C.c_uint32),
                ("reserved", C.c_uint32), ("payload", C.c_uint64)]


SBF, DBF, MLF = "SBF", "DBF", "MLF"     # fault kinds of relia_ntt_sim.cpp:109-132


def make_plan(entries):
    """entries: iterable of (limb, stage, butterfly, kind, payload) with kind 1 (xor) or 2 (replace)."""
    entries = list(entries)
    arr = (Inject * max(len(entries), 1))()
    for i, (limb, stage, bf, kind, payload) in enumerate(entries):
        arr[i] = Inject(limb, stage, bf, kind, 0, payload)
    return arr, len(entries)


def _nwt_forward_protected_host(self, a, L=None, start_mod=0, plan=()):
    """In-place protected forward NTT of a [n_poly][L][N]; returns (a, verdict[n_limbs], sums[n_limbs][2])."""
    L = self.L if L is None else L
    assert a.dtype == np.uint64 and a.flags.c_contiguous and a.size % (L * self.N) == 0
    n_poly = a.size // (L * self.N)
    verdict = np.zeros(n_poly * L, dtype=np.uint32)
    sums = np.zeros((n_poly * L, 2), dtype=np.uint64)
    arr, n = make_plan(plan)
    check(lib().b200fhe_nwt_forward_protected_host(self._h, _p(a), n_poly, L, start_mod,
                                                   verdict.ctypes.data_as(C.POINTER(C.c_uint32)), _p(sums),
                                                   C.cast(arr, vp), n))
    return a, verdict, sums


def _nwt_backward_protected_host(self, a, L=None, start_mod=0):
    """In-place protected inverse NTT of a [n_poly][L][N]; returns (a, verdict[n_limbs], sums[n_limbs][2]) with
    sums = (evaluation-side checksum of the inputs, coefficient-side checksum of the outputs)."""
    L = self.L if L is None else L
    assert a.dtype == np.uint64 and a.flags.c_contiguous and a.size % (L * self.N) == 0
    n_poly = a.size // (L * self.N)
    verdict = np.zeros(n_poly * L, dtype=np.uint32)
    sums = np.zeros((n_poly * L, 2), dtype=np.uint64)
    check(lib().b200fhe_nwt_backward_protected_host(self._h, _p(a), n_poly, L, start_mod,
                                                    verdict.ctypes.data_as(C.POINTER(C.c_uint32)), _p(sums)))
    return a, verdict, sums


def _abft_weights(self, idx):
    w = np.zeros(self.N, dtype=np.uint64); wh = np.zeros(self.N, dtype=np.uint64)
    check(lib().b200fhe_abft_weights(self._h, idx, _p(w), _p(wh)))
    return w, wh


def _pmult_accum_host(self, ct, pt, L=None, start_mod=0):
    """ct [T][P][L][N], pt [T][L][N] (numpy uint64) -> out [P][L][N] = sum_k ct_k * pt_k mod q."""
    L = self.L if L is None else L
    ct = np.ascontiguousarray(ct, dtype=np.uint64); pt = np.ascontiguousarray(pt, dtype=np.uint64)
    T, P = ct.shape[0], ct.shape[1]
    assert ct.shape == (T, P, L, self.N) and pt.shape == (T, L, self.N)
    out = np.zeros((P, L, self.N), dtype=np.uint64)
    check(lib().b200fhe_pmult_accum_host(self._h, _p(ct), _p(pt), T, _p(out), P, L, start_mod))
    return out


def _pmult_accum(self, ct_ptrs, pt_ptrs, out_ptr, n_poly, L=None, start_mod=0, accumulate=False, stream=0):
    """device pointers (ints): ct_ptrs[k] -> [P][L][N], pt_ptrs[k] -> [L][N], out_ptr -> [P][L][N]."""
    T = len(ct_ptrs)
    a = (vp * max(T, 1))(*ct_ptrs); b = (vp * max(T, 1))(*pt_ptrs)
    check(lib().b200fhe_pmult_accum(self._h, a, b, T, out_ptr, n_poly, self.L if L is None else L, start_mod,
                                    1 if accumulate else 0, stream))


Context.nwt_forward_protected_host = _nwt_forward_protected_host
Context.nwt_backward_protected_host = _nwt_backward_protected_host
Context.abft_weights = _abft_weights
Context.pmult_accum_host = _pmult_accum_host
Context.pmult_accum = _pmult_accum


class BaseConv:
    """ModUp plan (b200fhe_bconv): source base = context moduli src_idx, targets = context moduli dst_idx."""

    HPS, EXACT = 0, 1

    def __init__(self, ctx, src_idx, dst_idx, scale=1, mode=0):
        self.ctx, self.src, self.dst = ctx, list(src_idx), list(dst_idx)
        self.m, self.K, self.N = len(self.src), len(self.dst), ctx.N
        h = vp()
        if mode:
            assert scale == 1, "the exact-CRT mode has no scaled flavour"
            check(lib().b200fhe_bconv_create_mode(ctx._h, (i32 * self.m)(*self.src), self.m, (i32 * self.K)(*self.dst), self.K, mode, C.byref(h)))
        else:
            check(lib().b200fhe_bconv_create_scaled(ctx._h, (i32 * self.m)(*self.src), self.m, (i32 * self.K)(*self.dst), self.K, scale,
                                                    C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            lib().b200fhe_bconv_destroy(self._h)
            self._h = None

    __del__ = close

    def scale(self, x_ptr, y_ptr, j0, count, stream=0):
        check(lib().b200fhe_bconv_scale(self._h, x_ptr, y_ptr, j0, count, stream))

    def contract(self, y_ptr, out_ptr, k0, kcount, stream=0):
        check(lib().b200fhe_bconv_contract(self._h, y_ptr, out_ptr, k0, kcount, stream))

    def modup(self, in_ptr, out_ptr, stream=0):
        check(lib().b200fhe_bconv_modup(self._h, in_ptr, out_ptr, stream))

    def modup_static(self, in_ptr, out_ptr, stream=0):
        """b200fhe_bconv_modup_static: same conversion, scratch in the library's static device array (no allocation)."""
        check(lib().b200fhe_bconv_modup_static(self._h, in_ptr, out_ptr, stream))

    def modup_host(self, x):
        """x [m][N] uint64 residues mod the source moduli -> [K][N] residues mod the targets."""
        x = np.ascontiguousarray(x, dtype=np.uint64)
        assert x.shape == (self.m, self.N)
        out = np.zeros((self.K, self.N), dtype=np.uint64)
        check(lib().b200fhe_bconv_modup_host(self._h, _p(x), _p(out)))
        return out


class BconvInject(C.Structure):
    """b200fhe_bconv_inject: stage 1 = scale product, 2 = contraction output, 3 = y in memory between the kernels."""
    _fields_ = [("stage", C.c_uint32), ("limb", C.c_uint32), ("coeff", C.c_uint64), ("xor_mask", C.c_uint64)]


def _modup_protected_host(self, x, inject=()):
    """Checksum-carrying ModUp (baseConv.py:42-98 bConv_ecc on the device): x [m][N] -> (out [K][N], verdict [m+K],
    sums [m+K][2]).  inject: (stage, limb, coeff, xor_mask) tuples."""
    x = np.ascontiguousarray(x, dtype=np.uint64)
    assert x.shape == (self.m, self.N)
    out = np.zeros((self.K, self.N), dtype=np.uint64)
    verdict = np.zeros(self.m + self.K, dtype=np.uint32)
    sums = np.zeros((self.m + self.K, 2), dtype=np.uint64)
    arr = (BconvInject * max(len(inject), 1))()
    for i, e in enumerate(inject):
        arr[i] = BconvInject(*[int(v) for v in e])
    check(lib().b200fhe_bconv_modup_protected_host(self._h, _p(x), _p(out), verdict.ctypes.data_as(C.POINTER(C.c_uint32)), _p(sums),
                                                   C.cast(arr, vp), len(inject)))
    return out, verdict, sums


def _modup_protected(self, in_ptr, out_ptr, verdict_ptr, sums_ptr=None, stream=0):
    check(lib().b200fhe_bconv_modup_protected(self._h, in_ptr, out_ptr, verdict_ptr, sums_ptr, None, 0, stream))


BaseConv.modup_protected_host = _modup_protected_host
BaseConv.modup_protected = _modup_protected


class KeySwitch:
    """Hybrid key-switch plan (b200fhe_keyswitch): ctx holds n_q ciphertext moduli then n_p special primes.
    Device pointers in and out (torch tensors' data_ptr()); shapes in include/b200fhe.h."""

    def __init__(self, ctx, n_q, n_p, dnum, plain_modulus=1):
        self.ctx, self.n_q, self.n_p, self.dnum, self.N = ctx, n_q, n_p, dnum, ctx.N
        h = vp()
        check(lib().b200fhe_keyswitch_create_bgv(ctx._h, n_q, n_p, dnum, plain_modulus, C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            lib().b200fhe_keyswitch_destroy(self._h)
            self._h = None

    __del__ = close

    def workspace_bytes(self):
        w = sz()
        check(lib().b200fhe_keyswitch_info(self._h, None, None, None, C.byref(w)))
        return int(w.value)

    def modup(self, c_ptr, up_ptr, stream=0):
        check(lib().b200fhe_keyswitch_modup(self._h, c_ptr, up_ptr, stream))

    def moddown(self, acc_q_ptr, acc_p_ptr, out_ptr, n_poly, stream=0):
        check(lib().b200fhe_keyswitch_moddown(self._h, acc_q_ptr, acc_p_ptr, out_ptr, n_poly, stream))

    def apply(self, c_ptr, evk_q_ptr, evk_p_ptr, out_ptr, stream=0):
        check(lib().b200fhe_keyswitch_apply(self._h, c_ptr, evk_q_ptr, evk_p_ptr, out_ptr, stream))

    def set_inject(self, conversion, inject):
        """arm (stage, limb, coeff, xor_mask) faults for the next apply_protected; conversion: digit index, or dnum / dnum+1 for ModDown"""
        arr = (BconvInject * max(len(inject), 1))()
        for i, e in enumerate(inject):
            arr[i] = BconvInject(*[int(v) for v in e])
        check(lib().b200fhe_keyswitch_set_inject(self._h, conversion, C.cast(arr, vp), len(inject)))

    def apply_protected(self, c_ptr, evk_q_ptr, evk_p_ptr, out_ptr, verdict_ptr, stream=0):
        """verdict: uint32 [dnum + 2][n_q + n_p] on the device; all zero = every conversion check passed."""
        check(lib().b200fhe_keyswitch_apply_protected(self._h, c_ptr, evk_q_ptr, evk_p_ptr, out_ptr, verdict_ptr, stream))


def _ks_ntt_count(self):
    n = sz()
    check(lib().b200fhe_keyswitch_ntt_count(self._h, C.byref(n)))
    return int(n.value)


def _ks_apply_protected_full(self, c_ptr, evk_q_ptr, evk_p_ptr, out_ptr, verdict_ptr, verdict_ntt_ptr, stream=0):
    """apply_protected with every limb-transform protected too; verdict_ntt: uint32 [ntt_count()] on the device."""
    check(lib().b200fhe_keyswitch_apply_protected_full(self._h, c_ptr, evk_q_ptr, evk_p_ptr, out_ptr, verdict_ptr, verdict_ntt_ptr, stream))


KeySwitch.ntt_count = _ks_ntt_count
KeySwitch.apply_protected_full = _ks_apply_protected_full


def flip_bit(ptr, idx, bit, stream=0):
    check(lib().b200fhe_flip_bit(ptr, idx, bit, stream))


class FsInject(C.Structure):
    """b200fhe_fs_inject: `vector[row] += delta` after matrix stage 1 or 2 (four_step_ntt_prot.py:271-274)."""
    _fields_ = [("transform", C.c_uint32), ("stage", C.c_uint32), ("row", C.c_uint64), ("delta", C.c_uint64)]


def four_step_protected_host(a, mod, g, plan=(), device=0):
    """a [batch][N] uint64 -> (y [batch][N], checks [batch][4] = lhs1, rhs1, lhs2, rhs2, verdict [batch])."""
    a = np.ascontiguousarray(a, dtype=np.uint64)
    batch, N = a.shape
    y = np.zeros_like(a)
    checks = np.zeros((batch, 4), dtype=np.uint64)
    verdict = np.zeros(batch, dtype=np.uint32)
    plan = list(plan)
    arr = (FsInject * max(len(plan), 1))()
    for i, (tr, stage, row, delta) in enumerate(plan):
        arr[i] = FsInject(tr, stage, row, delta)
    check(lib().b200fhe_four_step_protected_host(mod, g, N, _p(a), _p(y), batch, _p(checks),
                                                 verdict.ctypes.data_as(C.POINTER(C.c_uint32)), C.cast(arr, vp), len(plan), device))
    return y, checks, verdict


class ReliaFault(C.Structure):
    """b200fhe_relia_fault: one corrupted operation of one trial (relia_ntt_sim.cpp's running op index)."""
    _fields_ = [("trial", C.c_uint32), ("kind", C.c_uint32), ("op", C.c_uint64), ("payload", C.c_uint64)]


def relia_op_ranges(N):
    """(s1_end, s2_end, s3_end): the simulator's OpRanges for N = n1 n2."""
    a, b, c = u64(), u64(), u64()
    check(lib().b200fhe_relia_op_ranges(N, C.byref(a), C.byref(b), C.byref(c)))
    return int(a.value), int(b.value), int(c.value)


def relia_four_step_host(q, g, a, weights, faults=(), fold_bits=4, want_out=True, device=0):
    """four_step_ntt of relia_ntt_sim.cpp for a [trials][N]; weights [trials][2 n2 + n1]; faults: iterable of
    (trial, op, kind, payload) or a ready (ReliaFault * n) array.  Returns (out or None, flags [trials] uint32)."""
    a = np.ascontiguousarray(a, dtype=np.uint64)
    weights = np.ascontiguousarray(weights, dtype=np.uint64)
    trials, N = a.shape
    if isinstance(faults, np.ndarray):          # structured array with b200fhe_relia_fault's layout (24-byte records)
        assert faults.dtype.itemsize == C.sizeof(ReliaFault) and faults.flags.c_contiguous
        arr, nf = faults.ctypes.data_as(vp), len(faults)
    elif isinstance(faults, C.Array):
        arr, nf = faults, len(faults)
    else:
        faults = list(faults)
        nf = len(faults)
        arr = (ReliaFault * max(nf, 1))()
        for i, (tr, op, kind, payload) in enumerate(faults):
            arr[i] = ReliaFault(int(tr), int(kind), int(op), int(payload))
    out = np.zeros_like(a) if want_out else None
    flags = np.zeros(trials, dtype=np.uint32)
    check(lib().b200fhe_relia_four_step_host(q, g, N, fold_bits, _p(a), _p(out) if want_out else None, _p(weights), trials,
                                             C.cast(arr, vp), nf, flags.ctypes.data_as(C.POINTER(C.c_uint32)), device))
    return out, flags


def _relia_faults(faults):
    if isinstance(faults, np.ndarray):          # structured array with b200fhe_relia_fault's layout (24-byte records)
        assert faults.dtype.itemsize == C.sizeof(ReliaFault) and faults.flags.c_contiguous
        return faults.ctypes.data_as(vp), len(faults)
    if isinstance(faults, C.Array):
        return faults, len(faults)
    faults = list(faults)
    arr = (ReliaFault * max(len(faults), 1))()
    for i, (tr, op, kind, payload) in enumerate(faults):
        arr[i] = ReliaFault(int(tr), int(kind), int(op), int(payload))
    return arr, len(faults)


def relia_standard_host(q, g, a, weights, faults=(), want_out=True, device=0):
    """One trial per row of non_four_step_ntt_sim.cpp (standard NTT + lazy bucket check + global check): a [trials][N],
    weights [trials][N]; faults as relia_four_step_host, ops in [0, (N/2) log2 N).  Returns (out or None, flags): bit 0 lazy,
    bit 1 global."""
    a = np.ascontiguousarray(a, dtype=np.uint64)
    weights = np.ascontiguousarray(weights, dtype=np.uint64)
    trials, N = a.shape
    assert weights.shape == a.shape
    arr, nf = _relia_faults(faults)
    out = np.zeros_like(a) if want_out else None
    flags = np.zeros(trials, dtype=np.uint32)
    check(lib().b200fhe_relia_standard_host(q, g, N, _p(a), _p(out) if want_out else None, _p(weights), trials, C.cast(arr, vp), nf,
                                            flags.ctypes.data_as(C.POINTER(C.c_uint32)), device))
    return out, flags


class FoldFault(C.Structure):
    """b200fhe_fold_fault: a fault at one of the Barrett multiplier's two product outputs (barrett_final.cpp:6-7)."""
    _fields_ = [("index", C.c_uint64), ("payload", C.c_uint64), ("product", C.c_uint32), ("kind", C.c_uint32)]


def barrett_params(q):
    K, mu = i32(), u64()
    check(lib().b200fhe_barrett_params(q, C.byref(K), C.byref(mu)))
    return int(K.value), int(mu.value)


def barrett_fold_host(q, a, b, fold, k=None, faults=(), propagate=0, device=0):
    """f4: element-wise Barrett products of a, b [n_vec][vec_len] on the device with the per-group XOR signatures
    (barrett_final.cpp "instra" / "sum") and, when k is given, the mod 2^k + 1 residue verdicts (outer_check.cpp
    "intra" bit 0 / "inter" bit 1, subtractor bit 2).  faults: (index, product, kind, payload) or a FoldFault array.
    Returns (c, sig_intra, sig_sum, verdict or None)."""
    a = np.ascontiguousarray(a, dtype=np.uint64); b = np.ascontiguousarray(b, dtype=np.uint64)
    n_vec, vec_len = a.shape
    groups = n_vec * ((vec_len + fold - 1) // fold)
    c = np.zeros_like(a)
    s1, s2 = np.zeros(groups, dtype=np.uint64), np.zeros(groups, dtype=np.uint64)
    v = np.zeros(groups, dtype=np.uint32)
    if isinstance(faults, np.ndarray):
        assert faults.dtype.itemsize == C.sizeof(FoldFault) and faults.flags.c_contiguous
        arr, nf = faults.ctypes.data_as(vp), len(faults)
    else:
        faults = list(faults)
        nf = len(faults)
        arr = (FoldFault * max(nf, 1))()
        for i, (idx, prod, kind, payload) in enumerate(faults):
            arr[i] = FoldFault(int(idx), int(payload), int(prod), int(kind))
        arr = C.cast(arr, vp)
    check(lib().b200fhe_barrett_fold_host(q, _p(a), _p(b), _p(c), n_vec, vec_len, fold, k or 0, 1 if k else 0, propagate, arr, nf,
                                          _p(s1), _p(s2), v.ctypes.data_as(C.POINTER(C.c_uint32)), device))
    return c, s1, s2, (v if k else None)


def modup_sharded_layout(n_q, n_p, dnum, world, rank):
    """(own_limbs, source_order [dnum][alpha], counts [dnum][world]) of b200fhe_modup_sharded_layout (pure host)."""
    alpha, total = n_q // dnum, n_q + n_p
    nq, nt = i32(), i32()
    own = (i32 * total)(); src = (i32 * (dnum * alpha))(); cnt = (i32 * (dnum * world))()
    check(lib().b200fhe_modup_sharded_layout(n_q, n_p, dnum, world, rank, C.byref(nq), C.byref(nt), own, src, cnt))
    return (list(own[:nt.value]), [list(src[d * alpha:(d + 1) * alpha]) for d in range(dnum)],
            [list(cnt[d * world:(d + 1) * world]) for d in range(dnum)])


class Shard:
    """b200fhe_shard: this rank's place in the world (+ its NCCL communicator when world > 1).  `dist` = an initialised
    torch.distributed (any backend) used only to hand rank 0's NCCL unique id to the other ranks."""

    def __init__(self, rank=0, world=1, device=0, dist=None):
        self.rank, self.world, self.device = rank, world, device
        idb = (C.c_uint8 * 128)()
        if world > 1 and dist is None:          # own transport: scale / raise halves only (b200fhe_shard_init with a NULL communicator)
            import torch
            torch.cuda.set_device(device)
            h = vp()
            check(lib().b200fhe_shard_init(None, rank, world, C.byref(h)))
            self._h = h
            return
        if world > 1:
            import torch
            if rank == 0:
                check(lib().b200fhe_shard_unique_id(idb))
            t = torch.tensor(list(idb), dtype=torch.uint8)
            if dist.get_backend() == "nccl":
                t = t.cuda(device)
            dist.broadcast(t, src=0)
            idb = (C.c_uint8 * 128)(*[int(v) for v in t.cpu().tolist()])
        h = vp()
        check(lib().b200fhe_shard_init_id(idb, rank, world, device, C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            lib().b200fhe_shard_destroy(self._h)
            self._h = None

    __del__ = close


class ModUpSharded:
    """b200fhe_modup_sharded plan of one rank."""

    def __init__(self, shard, log_n, moduli, n_q, n_p, dnum):
        self.shard, self.N, self.n_q, self.n_p, self.dnum = shard, 1 << log_n, n_q, n_p, dnum
        arr = np.array([int(q) for q in moduli], dtype=np.uint64)
        h = vp()
        check(lib().b200fhe_modup_sharded_create(shard._h, log_n, _p(arr), n_q, n_p, dnum, C.byref(h)))
        self._h = h
        nq, nt = i32(), i32()
        own = (i32 * (n_q + n_p))()
        check(lib().b200fhe_modup_sharded_info(h, C.byref(nq), C.byref(nt), own))
        self.n_own_q, self.n_own_total, self.own_limbs = nq.value, nt.value, list(own[:nt.value])

    def run(self, d_c_own, d_up_own, stream=None):
        check(lib().b200fhe_modup_sharded(self._h, d_c_own, d_up_own, stream))

    def scale(self, d_c_own, d_y_own, stream=None):
        check(lib().b200fhe_modup_sharded_scale(self._h, d_c_own, d_y_own, stream))

    def raise_(self, d_c_own, y_full_ptrs, d_up_own, stream=None):
        arr = (vp * len(y_full_ptrs))(*y_full_ptrs)
        check(lib().b200fhe_modup_sharded_raise(self._h, d_c_own, arr, d_up_own, stream))

    def close(self):
        if getattr(self, "_h", None):
            lib().b200fhe_modup_sharded_destroy(self._h)
            self._h = None

    __del__ = close


class Bgv:
    """b200fhe_bgv: the BGV host roles + evaluator composition of dotprod_test.cu (context, keys, encoder, encrypt / decrypt,
    multiply, relinearize, mod_switch_to_next, rotate, add).  Ciphertexts are device pointers [size][n_limbs][N]."""

    def __init__(self, log_n, coeff_moduli, special_size, plain_modulus, seed=1, device=0):
        arr = np.array([int(q) for q in coeff_moduli], dtype=np.uint64)
        h = vp()
        check(lib().b200fhe_bgv_create(log_n, _p(arr), len(arr), special_size, plain_modulus, seed, device, C.byref(h)))
        self._h, self.N, self.t = h, 1 << log_n, int(plain_modulus)
        self.n_q, self.n_p = len(arr) - special_size, special_size
        check(lib().b200fhe_bgv_gen_secretkey(h))
        check(lib().b200fhe_bgv_gen_publickey(h))

    def encode(self, values):
        v = np.ascontiguousarray(values, dtype=np.uint64)
        out = np.zeros(self.N, dtype=np.uint64)
        check(lib().b200fhe_bgv_encode(self._h, _p(v), v.size, _p(out)))
        return out

    def decode(self, plain):
        out = np.zeros(self.N, dtype=np.uint64)
        check(lib().b200fhe_bgv_decode(self._h, _p(np.ascontiguousarray(plain, dtype=np.uint64)), _p(out)))
        return out

    def encrypt(self, plain, d_ct):
        check(lib().b200fhe_bgv_encrypt(self._h, _p(np.ascontiguousarray(plain, dtype=np.uint64)), d_ct))

    def decrypt(self, d_ct, size, n_limbs, correction=1):
        out = np.zeros(self.N, dtype=np.uint64)
        check(lib().b200fhe_bgv_decrypt(self._h, d_ct, size, n_limbs, correction, _p(out)))
        return out

    def multiply(self, d_a, d_b, d_out, n_limbs):
        check(lib().b200fhe_bgv_multiply(self._h, d_a, d_b, d_out, n_limbs))

    def relinearize(self, d_ct, n_limbs):
        check(lib().b200fhe_bgv_relinearize(self._h, d_ct, n_limbs))

    def mod_switch_to_next(self, d_ct, size, n_limbs, d_out, correction=1):
        cf = u64(correction)
        check(lib().b200fhe_bgv_mod_switch_to_next(self._h, d_ct, size, n_limbs, d_out, C.byref(cf)))
        return int(cf.value)

    def rotate(self, d_ct, n_limbs, step, columns=False):
        check(lib().b200fhe_bgv_rotate(self._h, d_ct, n_limbs, step, 1 if columns else 0))

    def add(self, d_a, d_b, size, n_limbs):
        check(lib().b200fhe_bgv_add(self._h, d_a, d_b, size, n_limbs))

    def close(self):
        if getattr(self, "_h", None):
            lib().b200fhe_bgv_destroy(self._h)
            self._h = None

    __del__ = close


def fault_payload(subtype, bitwidth, rng):
    """XOR mask of one fault of the reference's taxonomy (rfhe_framewk/src/four_step_ntt_protected.cpp:163-245), for
    b200fhe_inject kind 1 (computation fault on a butterfly product) or kind 3 (storage fault between the passes):
      SCF-BF   one random bit below `bitwidth`                                   (:170-172, :186-188)
      SCF-MBU  a burst of K adjacent bits, K in {2, 3, 4} (vectors: {2, 3})      (:173-178, :189-194)
      MCF-PPE  one random bit (partial-product error)                            (:203-206, :233-235)
      MCF-CTE  one bit of the upper half (carry-tree error)                      (:207-211)
    MCF-CLE (a whole output line shifted by a multiple of one operand product, :212-222) belongs to the matrix-multiply
    form of the prototype's stages and has no butterfly counterpart; it is not generated here.
    rng: numpy Generator."""
    if subtype in ("SCF-BF", "MCF-PPE"):
        return 1 << int(rng.integers(0, bitwidth))
    if subtype == "SCF-MBU":
        K = int(rng.choice([2, 3, 4])) if bitwidth >= 4 else 2
        start = int(rng.integers(0, max(0, bitwidth - K) + 1))
        return ((1 << K) - 1) << start
    if subtype == "MCF-CTE":
        lo = max(1, bitwidth // 2)
        return 1 << int(rng.integers(lo, max(lo, bitwidth - 1) + 1))
    raise ValueError("unknown fault subtype %r" % (subtype,))
