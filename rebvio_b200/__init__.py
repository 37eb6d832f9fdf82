"""rebvio_b200 — B200-native (sm_100a) edge front end + edge tracker of REBVIO behind a C ABI.

The product is `librebvio_b200.so` (hand-written CUDA, rebvio_b200/csrc) plus the C++ adapter headers in
rebvio_b200/cpp that carry the reference's class API. This Python package only binds the C ABI with ctypes
(capi.py), provides the synthetic sequence generator (synth.py) and a thin host pipeline used by the tests and
bench.py (pipeline.py). There is no CPU fallback: importing `capi` without the built library raises.
"""
__version__ = "0.1.0"
