"""ctypes binding of libzangcu.so (include/zangcu.h) and the numpy dtypes of its PODs.

The library is the product; this file only declares its C ABI for the Python harness (tests, bench).
There is no fallback: if the shared library is missing the import fails loudly.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ZANGCU_LIB", os.path.join(_HERE, "libzangcu.so"))  # override: A/B builds of the same ABI

from .abi import *  # noqa: F401,F403  (enums, dtypes and PODs of zangcu.h; pure Python)
from .abi import ptr  # noqa: F401

# every exported symbol of include/zangcu.h with (restype, argtypes); pointers are void* on purpose
_P, _I, _U, _F, _S, _U64 = C.c_void_p, C.c_int, C.c_uint32, C.c_float, C.c_size_t, C.c_uint64
PROTOTYPES = {
    "zcu_abi_version": (_I, []),
    "zcu_device_count": (_I, []),
    "zcu_ctx_create": (_I, [_I, _P]),
    "zcu_ctx_create_on_stream": (_I, [_I, _P, _P]),
    "zcu_ctx_destroy": (None, [_P]),
    "zcu_ctx_sync": (_I, [_P]),
    "zcu_last_error": (C.c_char_p, [_P]),
    "zcu_ctx_stream": (_P, [_P]),
    "zcu_ctx_launch_count": (_U64, [_P]),
    "zcu_ctx_set_paint_mode": (_I, [_P, _I]),
    "zcu_ctx_paint_mode": (_I, [_P]),
    "zcu_ctx_set_paint_output": (_I, [_P, _I]),
    "zcu_ctx_paint_output": (_I, [_P]),
    "zcu_ctx_noise_resync_count": (_I, [_P, _P]),
    "zcu_ctx_paint_warm_stats": (_I, [_P, _P]),
    "zcu_xoshiro_jump_host": (_I, [_P, _U]),
    "zcu_selftest_div_by_const": (_I, [_P, _F, _P, _P]),
    "zcu_selftest_div_numerator": (_I, [_P, _F, _P, _P]),
    "zcu_timer_start": (_I, [_P]),
    "zcu_timer_stop_ms": (_I, [_P, _P]),
    "zcu_buf_alloc": (_I, [_P, _U, _U, _P]),
    "zcu_buf_wrap": (_I, [_P, _P, _U, _U, _P]),
    "zcu_buf_free": (None, [_P]),
    "zcu_buf_upload": (_I, [_P, _P, _P]),
    "zcu_buf_download": (_I, [_P, _P, _P]),
    "zcu_buf_device_ptr": (_P, [_P]),
    "zcu_buf_frames": (_U, [_P]),
    "zcu_buf_voices": (_U, [_P]),
    "zcu_bank_create": (_I, [_P, _I, _U, _U, _P]),
    "zcu_bank_free": (None, [_P]),
    "zcu_bank_reset": (_I, [_P, _P, _U]),
    "zcu_bank_state_size": (_S, [_I]),
    "zcu_bank_get_state": (_I, [_P, _P, _P, _S]),
    "zcu_bank_set_state": (_I, [_P, _P, _P, _S]),
    "zcu_delay_bank_create": (_I, [_P, _I, _U, _U, _P]),
    "zcu_bank_delay_samples": (_U, [_P]),
    "zcu_bank_get_delay_line": (_I, [_P, _P, _P, _S]),
    "zcu_bank_set_delay_line": (_I, [_P, _P, _P, _S]),
    "zcu_spans_upload": (_I, [_P, _U, _U, _P, _P, _P]),
    "zcu_spans_full": (_I, [_P, _U, _U, _U, _P]),
    "zcu_spans_free": (None, [_P]),
    "zcu_zero": (_I, [_P, _P, _P]),
    "zcu_set": (_I, [_P, _P, _P, _F]),
    "zcu_copy": (_I, [_P, _P, _P, _P]),
    "zcu_add": (_I, [_P, _P, _P, _P, _P]),
    "zcu_add_into": (_I, [_P, _P, _P, _P]),
    "zcu_add_scalar": (_I, [_P, _P, _P, _P, _F]),
    "zcu_add_scalar_into": (_I, [_P, _P, _P, _F]),
    "zcu_multiply": (_I, [_P, _P, _P, _P, _P]),
    "zcu_multiply_with": (_I, [_P, _P, _P, _P]),
    "zcu_multiply_scalar": (_I, [_P, _P, _P, _P, _F]),
    "zcu_multiply_with_scalar": (_I, [_P, _P, _P, _F]),
    "zcu_set_rows": (_I, [_P, _P, _P, _P, _U]),
    "zcu_add_scalar_rows": (_I, [_P, _P, _P, _P, _P, _U]),
    "zcu_add_scalar_into_rows": (_I, [_P, _P, _P, _P, _U]),
    "zcu_multiply_scalar_rows": (_I, [_P, _P, _P, _P, _P, _U]),
    "zcu_multiply_with_scalar_rows": (_I, [_P, _P, _P, _P, _U]),
    "zcu_paint_sine_osc": (_I, [_P, _P, _P, _P, _U, _P, _P, _P]),
    "zcu_paint_pulse_osc": (_I, [_P, _P, _P, _P, _U, _P, _P]),
    "zcu_paint_trisaw_osc": (_I, [_P, _P, _P, _P, _U, _P, _P]),
    "zcu_paint_noise": (_I, [_P, _P, _P, _P, _U, _P]),
    "zcu_paint_envelope": (_I, [_P, _P, _P, _P, _U, _P]),
    "zcu_paint_gate": (_I, [_P, _P, _P, _U, _P]),
    "zcu_paint_filter": (_I, [_P, _P, _P, _P, _U, _P, _P, _P, _P]),
    "zcu_paint_portamento": (_I, [_P, _P, _P, _P, _U, _P]),
    "zcu_paint_curve": (_I, [_P, _P, _P, _P, _U, _P, _U, _P]),
    "zcu_paint_sampler": (_I, [_P, _P, _P, _P, _U, _P, _P]),
    "zcu_paint_decimator": (_I, [_P, _P, _P, _P, _U, _P, _P]),
    "zcu_paint_distortion": (_I, [_P, _P, _P, _U, _P, _P]),
    "zcu_paint_cycle": (_I, [_P, _P, _P, _P, _U, _P, _P]),
    "zcu_paint_fm_oscillator": (_I, [_P, _P, _P, _P, _U, _P, _P, _P]),
    "zcu_paint_simple_delay": (_I, [_P, _P, _P, _P, _P]),
    "zcu_paint_filtered_echoes": (_I, [_P, _P, _P, _P, _U, _P, _P]),
    "zcu_filter_cutoff_from_frequency": (_F, [_F, _F]),
    "zcu_sample_create": (_I, [_P, _P, _S, _U, _U, _I, _P]),
    "zcu_sample_free": (None, [_P]),
    "zcu_mix_reduce": (_I, [_P, _P, _P, _U, _P]),
    "zcu_mix_ordered": (_I, [_P, _P, _P, _U, _P]),
    "zcu_mix_down": (_I, [_P, _P, _I, _F, _P, _S]),
    "zcu_mix_down_static": (_I, [_P, _P, _I, _F, _P, _S]),
    "zcu_trigger_init": (None, [_P]),
    "zcu_trigger_reset": (None, [_P]),
    "zcu_trigger_run": (_U, [_P, _U, _U, _P, _P, _U, _U, _P, _U]),
    "zcu_impulse_queue_init": (None, [_P]),
    "zcu_impulse_queue_push": (None, [_P, _U, _U, _U]),
    "zcu_impulse_queue_consume": (_U, [_P, _P, _P]),
    "zcu_id_generator_init": (None, [_P]),
    "zcu_id_generator_next": (_U, [_P]),
    "zcu_note_tracker_init": (None, [_P, _P, _U]),
    "zcu_note_tracker_reset": (None, [_P]),
    "zcu_note_tracker_consume": (_U, [_P, _F, _U, _U, _P, _P, _U]),
    "zcu_poly_dispatcher_init": (_I, [_P, _U]),
    "zcu_poly_dispatcher_reset": (None, [_P]),
    "zcu_poly_dispatcher_dispatch": (_I, [_P, _P, _P, _P, _U, _P, _P, _P]),
    "zcu_ramp_advance": (_U, [_P, _F, _U, _P]),
    "zcu_poly_voice_init": (_I, [_P, _P, _U, _U]),
    "zcu_poly_voice_reset": (None, [_P]),
    "zcu_poly_voice_spans": (C.c_int64, [_P, _P, _F, _U, _U, _U, _P, _P, _U64]),
    "zcu_accum_advance": (None, [_P, _F, _U]),
    "zcu_cycle_advance": (None, [_P, _F, _U]),
    "zcu_bank_trigger_init": (None, [_P, _U, _P]),
    "zcu_bank_spans": (C.c_int64, [_P, _U, _P, _P, _U, _U, _U, _U, _P, _P, _U64]),
    "zcu_poly_bank_create": (_I, [_P, _U, _U, _P, _P, _P, _U, _F, _U, _P]),
    "zcu_poly_bank_free": (None, [_P]),
    "zcu_poly_bank_reset": (_I, [_P, _P]),
    "zcu_poly_bank_get_state": (_I, [_P, _P, _P]),
    "zcu_poly_bank_spans": (_I, [_P, _P, _U, _U, _P]),
    "zcu_event_bank_create": (_I, [_P, _U, _P, _P, _U, _U, _P]),
    "zcu_event_bank_free": (None, [_P]),
    "zcu_event_bank_reset": (_I, [_P, _P]),
    "zcu_event_bank_get_state": (_I, [_P, _P, _P]),
    "zcu_event_bank_set_state": (_I, [_P, _P, _P]),
    "zcu_event_bank_spans": (_I, [_P, _P, _U, _U, _P]),
    "zcu_spans_download": (_I, [_P, _P, _P, _P, _U64, _P]),
    "zcu_spans_prev_note_on_views": (_I, [_P, _P, _P, _P, _P, _P]),
    "zcu_render_create": (_I, [_P, _P, _P, _P, _P]),
    "zcu_render_free": (None, [_P]),
    "zcu_render_set_events": (_I, [_P, _P, _P, _P]),
    "zcu_render_run": (_I, [_P, _P, _P]),
    "zcu_render_run_voices": (_I, [_P, _P, _U, _P]),
    "zcu_render_run_voice_window": (_I, [_P, _P, _U, _U, _P]),
    "zcu_render_launches_per_run": (_U64, [_P]),
    "zcu_render_set_profiling": (_I, [_P, _I]),
    "zcu_render_kernel_ms": (_I, [_P, _P, _P, _P]),
    "zcu_render_aux_kernel_ms": (_I, [_P, _P, _P, _P]),
    "zcu_render_set_segments": (_I, [_P, _U]),
    "zcu_render_set_filter_contraction": (_I, [_P, C.c_int]),
    "zcu_render_segment_info": (_I, [_P, _P, _P, _P, _P, _P]),
    "zcu_graph_create": (_I, [_P, _P, _P, _U, _P, _P]),
    "zcu_graph_free": (None, [_P]),
    "zcu_graph_source": (C.c_char_p, [_P]),
    "zcu_graph_compile_check": (_I, [_P, _P, _U, _P, _P, _S, _P]),
    "zcu_graph_set_field": (_I, [_P, _P, _U, _P, _P, _S, _U]),
    "zcu_graph_set_curve_nodes": (_I, [_P, _P, _U, _P, _U]),
    "zcu_graph_set_sample": (_I, [_P, _P, _U, _P]),
    "zcu_graph_set_gains": (_I, [_P, _P, _P]),
    "zcu_graph_paint": (_I, [_P, _P, _P, _P, _P]),
    "zcu_graph_set_segments": (_I, [_P, _U]),
    "zcu_graph_segment_info": (_I, [_P, _P, _P, _P, _P, _P]),
    "zcu_graph_warm_retries": (_I, [_P, _P, _P]),
    "zcu_nccl_unique_id": (_I, [_P]),
    "zcu_nccl_init": (_I, [_P, _I, _I, _P]),
    "zcu_nccl_reduce_mix": (_I, [_P, _P, _I]),
    "zcu_nccl_finalize": (_I, [_P]),
}


def load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback for the CUDA path)")
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # raises AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    return lib


lib = load()
