"""stark_mlwe_b200 — B200-native (sm_100a) implementation of stark-mlwe's commitment hot path.

The product is the CUDA shared library `libsmw.so` behind the C ABI in include/smw.h; this package is the
host-side mirror of the reference's Rust API for that path (same names, argument meaning and error
behaviour), used by the parity tests and the benchmark.  There is no CPU fallback: importing `api`
without the built library, or creating a Context without a CUDA device, fails loudly.
"""
__all__ = ["api", "build"]
