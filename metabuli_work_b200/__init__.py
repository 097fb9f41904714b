"""metabuli_work_b200 — B200-native implementation of Metabuli's `classify` hot path.

The product is native: `libmetabuli_b200.so` (hand-written sm_100a CUDA behind the C ABI of
include/metabuli_b200.h), `libmbhost.so` / `metabuli` (C++ host).  This package is only the ctypes
binding the tests and bench.py use; it holds no compute and has no CPU fallback — importing
`metabuli_work_b200.binding` fails loudly when the native libraries are missing.
"""
