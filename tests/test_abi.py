"""The drop-in boundary: libzangcu.so loads, exports every symbol include/zangcu.h declares, and fails
loudly (never silently falls back to a CPU path) when there is no CUDA device."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from zang_b200 import _lib as L
from zang_b200._lib import lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    src = open(os.path.join(ROOT, "include", "zangcu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(zcu_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported():
    names = declared_functions()
    assert len(names) > 80
    out = subprocess.run(["nm", "-D", "--defined-only", L.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r" T (zcu_[a-z0-9_]+)", out))
    missing = [n for n in names if n not in exported]
    assert not missing, f"declared in zangcu.h but not exported: {missing}"


def test_python_prototypes_cover_the_header():
    names = set(declared_functions())
    assert names == set(L.PROTOTYPES), (sorted(names - set(L.PROTOTYPES)), sorted(set(L.PROTOTYPES) - names))


def test_header_compiles_as_c_and_pods_match_numpy():
    # the header is plain C (usable from Zig's @cImport); struct sizes match the numpy dtypes the harness uses
    prog = r'''
#include "zangcu.h"
#include <stdio.h>
int main(void) {
  printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\n", sizeof(zcu_span), sizeof(zcu_envelope_params), sizeof(zcu_envelope_state),
         sizeof(zcu_noise_state), sizeof(zcu_curve_state), sizeof(zcu_portamento_params), sizeof(zcu_note_event), sizeof(zcu_impulse),
         sizeof(zcu_impulse_queue), sizeof(zcu_poly_dispatcher), sizeof(zcu_render_config), sizeof(zcu_filter_params));
  return 0; }'''
    d = os.path.join(ROOT, "oracle", "_build")
    os.makedirs(d, exist_ok=True)
    cfile, exe = os.path.join(d, "abi_probe.c"), os.path.join(d, "abi_probe")
    open(cfile, "w").write(prog)
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), cfile, "-o", exe], check=True)
    sizes = [int(x) for x in subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()]
    want = [L.span_dtype.itemsize, L.param_dtypes[L.MOD_ENVELOPE].itemsize, L.state_dtypes[L.MOD_ENVELOPE].itemsize,
            L.state_dtypes[L.MOD_NOISE].itemsize, L.state_dtypes[L.MOD_CURVE].itemsize, L.param_dtypes[L.MOD_PORTAMENTO].itemsize,
            L.note_event_dtype.itemsize, L.impulse_dtype.itemsize, C.sizeof(L.ImpulseQueuePOD), C.sizeof(L.PolyDispatcherPOD),
            C.sizeof(L.RenderConfig), L.param_dtypes[L.MOD_FILTER].itemsize]
    assert sizes == want


def test_state_sizes():
    for kind, dt in L.state_dtypes.items():
        assert lib.zcu_bank_state_size(kind) == (0 if dt is None else dt.itemsize)


def test_host_scalar_cutoff_matches_oracle():
    import oracle
    for f in (10.0, 55.0, 440.0, 3520.0, 8000.0, 23999.0, 30000.0):
        for sr in (44100.0, 48000.0):
            a = np.float32(lib.zcu_filter_cutoff_from_frequency(f, sr))
            b = np.float32(oracle.orc.orc_cutoff_from_frequency(f, sr))
            assert a.view(np.uint32) == b.view(np.uint32)


def test_no_cpu_fallback_without_device():
    if lib.zcu_device_count() > 0:
        pytest.skip("a CUDA device is present")
    h = C.c_void_p()
    assert lib.zcu_ctx_create(0, C.byref(h)) == L.ERR_NO_DEVICE and not h.value
    from zang_b200.engine import Context, ZangCuError
    with pytest.raises(ZangCuError):
        Context(0)
    # every compute entry point refuses a null context instead of computing on the host
    assert lib.zcu_zero(None, None, None) == L.ERR_INVALID
    assert lib.zcu_render_run(None, None, None) == L.ERR_INVALID


def test_product_does_not_reference_the_oracle():
    # the oracle is test infrastructure: nothing under zang_b200/ or include/ may mention it
    bad = []
    for base in ("zang_b200", "include"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            if "build" in dp.split(os.sep):
                continue
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                    txt = open(os.path.join(dp, f), errors="ignore").read()
                    if re.search(r"zang_oracle|oracle_capi|import oracle|oracle/", txt):
                        bad.append(os.path.join(dp, f))
    assert not bad, bad
