"""rosella_b200 — B200-native implementation of Rosella's data-parallel hot path
(TNF counting + exact TNF / coverage kNN graphs) behind a C ABI (include/rosella_gpu.h).

The compute lives in librosella_gpu.so (hand-written sm_100a CUDA); this package only holds the
ctypes binding, the host-side mirror of the reference interface and synthetic-input generators.
There is no CPU fallback: importing `rosella_b200.gpu` works anywhere, creating a `Context`
requires the built library and an sm_100 device.
"""
__version__ = "0.1.0"
