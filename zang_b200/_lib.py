"""ctypes binding of libzangcu.so (include/zangcu.h) and the numpy dtypes of its PODs.

The library is the product; this file only declares its C ABI for the Python harness (tests, bench).
There is no fallback: if the shared library is missing the import fails loudly.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ZANGCU_LIB", os.path.join(_HERE, "libzangcu.so"))  # override: A/B builds of the same ABI

# ---- error codes / enums (zangcu.h) ---------------------------------------------------------------
OK, ERR_INVALID, ERR_NO_DEVICE, ERR_CUDA, ERR_OOM, ERR_UNSUPPORTED, ERR_NCCL = range(7)
(MOD_SINE_OSC, MOD_PULSE_OSC, MOD_TRISAW_OSC, MOD_NOISE, MOD_ENVELOPE, MOD_GATE, MOD_FILTER, MOD_PORTAMENTO,
 MOD_CURVE, MOD_SAMPLER, MOD_DECIMATOR, MOD_DISTORTION, MOD_CYCLE, MOD_SIMPLE_DELAY, MOD_FILTERED_ECHOES, MOD_FM_OSCILLATOR) = range(16)
SPAN_NOTE_ID_CHANGED = 1
ENV_IDLE, ENV_ATTACK, ENV_DECAY, ENV_SUSTAIN, ENV_RELEASE = range(5)
CURVE_INSTANTANEOUS, CURVE_LINEAR, CURVE_SQUARED, CURVE_CUBED = range(4)
NOISE_WHITE, NOISE_PINK = 0, 1
FILTER_BYPASS, FILTER_LOW_PASS, FILTER_BAND_PASS, FILTER_HIGH_PASS, FILTER_NOTCH, FILTER_ALL_PASS = range(6)
INTERP_LINEAR, INTERP_SMOOTHSTEP = 0, 1
FMT_UNSIGNED8, FMT_SIGNED16_LSB, FMT_SIGNED24_LSB, FMT_SIGNED32_LSB = range(4)
DIST_OVERDRIVE, DIST_CLIP = 0, 1
AUDIO_SIGNED8, AUDIO_SIGNED16_LSB = 0, 1
INSTR_NICE = 0
RENDER_EXACT, RENDER_SCAN, RENDER_SEGMENTS = 0, 1, 2
PAINT_EXACT, PAINT_TOLERANCE = 0, 1
MAX_POLYPHONY = 32

# ---- POD dtypes -------------------------------------------------------------------------------------
span_dtype = np.dtype([("start", "<u4"), ("end", "<u4"), ("flags", "<u4"), ("param", "<u4")])
painter_state = [("t", "<f4"), ("last_value", "<f4"), ("start", "<f4")]
painter_curve = np.dtype([("kind", "<i4"), ("duration", "<f4")])

state_dtypes = {
    MOD_SINE_OSC: np.dtype([("t", "<f4")]),
    MOD_PULSE_OSC: np.dtype([("cnt", "<u4")]),
    MOD_TRISAW_OSC: np.dtype([("cnt", "<u4"), ("t", "<f4")]),
    MOD_NOISE: np.dtype([("s", "<u8", (4,))]),
    MOD_ENVELOPE: np.dtype([("state", "<i4"), ("painter", painter_state)]),
    MOD_FILTER: np.dtype([("l", "<f4"), ("b", "<f4")]),
    MOD_PORTAMENTO: np.dtype([("painter", painter_state)]),
    MOD_CURVE: np.dtype([("t", "<f4"), ("current_song_note", "<u4"), ("current_song_note_offset", "<i4"), ("next_song_note", "<u4")]),
    MOD_SAMPLER: np.dtype([("t", "<f4")]),
    MOD_DECIMATOR: np.dtype([("dval", "<f4"), ("dcount", "<f4")]),
    MOD_CYCLE: np.dtype([("t", "<f4")]),
    MOD_FM_OSCILLATOR: np.dtype([("t", "<f4"), ("feedback1", "<f4"), ("feedback2", "<f4")]),
    MOD_SIMPLE_DELAY: np.dtype([("index", "<u4")]),
    MOD_FILTERED_ECHOES: np.dtype([("index", "<u4"), ("l", "<f4"), ("b", "<f4")]),
    MOD_GATE: None,
    MOD_DISTORTION: None,
}

param_dtypes = {
    MOD_SINE_OSC: np.dtype([("sample_rate", "<f4"), ("freq", "<f4"), ("phase", "<f4")]),
    MOD_PULSE_OSC: np.dtype([("sample_rate", "<f4"), ("freq", "<f4"), ("color", "<f4")]),
    MOD_TRISAW_OSC: np.dtype([("sample_rate", "<f4"), ("freq", "<f4"), ("color", "<f4")]),
    MOD_NOISE: np.dtype([("color", "<i4")]),
    MOD_ENVELOPE: np.dtype([("sample_rate", "<f4"), ("attack", painter_curve), ("decay", painter_curve), ("release", painter_curve),
                            ("sustain_volume", "<f4"), ("note_on", "<i4")]),
    MOD_GATE: np.dtype([("note_on", "<i4")]),
    MOD_FILTER: np.dtype([("type", "<i4"), ("cutoff", "<f4"), ("res", "<f4")]),
    MOD_PORTAMENTO: np.dtype([("sample_rate", "<f4"), ("curve", painter_curve), ("goal", "<f4"), ("note_on", "<i4"), ("prev_note_on", "<i4")]),
    MOD_CURVE: np.dtype([("sample_rate", "<f4"), ("function", "<i4"), ("node_start", "<u4"), ("node_count", "<u4")]),
    MOD_SAMPLER: np.dtype([("sample_rate", "<f4"), ("channel", "<u4"), ("loop", "<i4")]),
    MOD_DECIMATOR: np.dtype([("sample_rate", "<f4"), ("fake_sample_rate", "<f4")]),
    MOD_DISTORTION: np.dtype([("type", "<i4"), ("ingain", "<f4"), ("outgain", "<f4"), ("offset", "<f4")]),
    MOD_CYCLE: np.dtype([("sample_rate", "<f4"), ("speed", "<f4")]),
    MOD_FM_OSCILLATOR: np.dtype([("sample_rate", "<f4"), ("waveform", "<u4"), ("feedback", "<f4")]),
    MOD_SIMPLE_DELAY: np.dtype([("unused", "<u4")]),
    MOD_FILTERED_ECHOES: np.dtype([("feedback_volume", "<f4"), ("cutoff", "<f4")]),
}
curve_node_dtype = np.dtype([("value", "<f4"), ("t", "<f4")])
impulse_dtype = np.dtype([("frame", "<u4"), ("note_id", "<u4"), ("event_id", "<u4")])
song_event_dtype = np.dtype([("t", "<f4"), ("note_id", "<u4"), ("param", "<u4")])
note_event_dtype = np.dtype([("frame", "<u4"), ("note_id", "<u4"), ("freq", "<f4"), ("note_on", "<i4")])
bank_trigger_dtype = np.dtype([("has_note", "<i4"), ("note_id", "<u4"), ("param", "<u4"), ("next_event", "<u4")])
poly_slot_dtype = np.dtype([("used", "<i4"), ("note_id", "<u4"), ("event_id", "<u4"), ("note_on", "<i4")])


class TriggerPOD(C.Structure):
    _fields_ = [("has_note", C.c_int32), ("note_id", C.c_uint32), ("param", C.c_uint32)]


class ImpulsePOD(C.Structure):
    _fields_ = [("frame", C.c_uint32), ("note_id", C.c_uint32), ("event_id", C.c_uint32)]


class ImpulseQueuePOD(C.Structure):
    _fields_ = [("impulses", ImpulsePOD * 32), ("params", C.c_uint32 * 32), ("length", C.c_uint32), ("next_event_id", C.c_uint32)]


class IdGeneratorPOD(C.Structure):
    _fields_ = [("next_id", C.c_uint32)]


class NoteTrackerPOD(C.Structure):
    _fields_ = [("song", C.c_void_p), ("song_len", C.c_uint32), ("next_song_event", C.c_uint32), ("t", C.c_float)]


class PolySlotPOD(C.Structure):
    _fields_ = [("used", C.c_int32), ("note_id", C.c_uint32), ("event_id", C.c_uint32), ("note_on", C.c_int32)]


class PolyDispatcherPOD(C.Structure):
    _fields_ = [("polyphony", C.c_uint32), ("slots", PolySlotPOD * MAX_POLYPHONY)]


class PolyVoicePOD(C.Structure):
    _fields_ = [("tracker", NoteTrackerPOD), ("dispatcher", PolyDispatcherPOD), ("triggers", TriggerPOD * MAX_POLYPHONY)]


class RenderConfig(C.Structure):
    _fields_ = [("instrument", C.c_int32), ("voices", C.c_uint32), ("buffer_frames", C.c_uint32), ("total_frames", C.c_uint32),
                ("block_buffers", C.c_uint32), ("channels", C.c_uint32), ("sample_rate", C.c_float), ("first_seed", C.c_uint32),
                ("mode", C.c_int32)]


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
    "zcu_ctx_noise_resync_count": (_I, [_P, _P]),
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
    "zcu_render_create": (_I, [_P, _P, _P, _P, _P]),
    "zcu_render_free": (None, [_P]),
    "zcu_render_set_events": (_I, [_P, _P, _P, _P]),
    "zcu_render_run": (_I, [_P, _P, _P]),
    "zcu_render_run_voices": (_I, [_P, _P, _U, _P]),
    "zcu_render_launches_per_run": (_U64, [_P]),
    "zcu_render_set_profiling": (_I, [_P, _I]),
    "zcu_render_kernel_ms": (_I, [_P, _P, _P, _P]),
    "zcu_render_aux_kernel_ms": (_I, [_P, _P, _P, _P]),
    "zcu_render_segment_info": (_I, [_P, _P, _P, _P, _P, _P]),
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


def ptr(a):
    """void* of a numpy array (or None)."""
    if a is None:
        return None
    return a.ctypes.data_as(C.c_void_p)
