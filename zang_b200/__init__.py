"""zang_b200 — B200-native batch synthesis engine behind zang's module contract.

The product is libzangcu.so (CUDA, sm_100a) with the C ABI of include/zangcu.h.  This package is the
Python harness over that ABI: `engine` mirrors the module contract, `events` the Notes/Trigger layer,
`banks` / `song` compose instrument graphs; all of them bind the library at import (`_lib`) and fail loudly
when it has not been built — there is no CPU fallback.  Only `abi` (dtypes / enums of zangcu.h) and the
workload generators in `synth` are pure Python, so that the CPU reference arm of bench.py can build the same
synthetic workload without mapping the CUDA library.
"""


def __getattr__(name):
    if name == "lib":
        from . import _lib
        return _lib.lib
    raise AttributeError(name)
