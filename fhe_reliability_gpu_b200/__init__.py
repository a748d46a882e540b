"""B200-native hot path of Stardust-lf/fhe-reliability-gpu: batched negacyclic
NTT/INTT over RNS limbs, protected four-step NTT, pmult, ModUp base conversion.
CUDA kernels + C ABI live in csrc/ (built into lib/libb200fhe.so); this package
is the ctypes mirror of the reference's harness/prototype interfaces."""
from .binding import B200FHEError, BaseConv, Context, KeySwitch, LIB_PATH, flip_bit, host_tables, lib, min_primitive_root, primes_create

__all__ = ["B200FHEError", "BaseConv", "Context", "KeySwitch", "LIB_PATH", "flip_bit", "host_tables", "lib", "min_primitive_root",
           "primes_create"]
