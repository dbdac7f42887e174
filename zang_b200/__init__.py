"""zang_b200 — B200-native batch synthesis engine behind zang's module contract.

The product is libzangcu.so (CUDA, sm_100a) with the C ABI of include/zangcu.h.  This package is the
Python harness over that ABI: `engine` mirrors the module contract, `events` the Notes/Trigger layer,
`synth` generates the synthetic workloads of the benchmark configurations.
"""
from . import _lib  # noqa: F401  (fails loudly when libzangcu.so has not been built)
from ._lib import lib  # noqa: F401
