"""Pure-Python mirror of the PODs and enums of include/zangcu.h: numpy dtypes, ctypes structures, constants.

Nothing here loads libzangcu.so.  The ctypes binding (`_lib`) re-exports these names; the test-side oracle loader
(tests/oracle.py) and the workload generators (`synth`) import them from here, so the CPU reference arm of bench.py
runs without mapping the CUDA library.
"""
import ctypes as C

import numpy as np

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
OUT_ACCUMULATE, OUT_OVERWRITE = 0, 1
MAX_POLYPHONY = 32
# fused module graphs
GREF_NONE, GREF_TEMP, GREF_OUTPUT, GREF_INPUT = range(4)
GI_BASIC, GI_ARITH, GI_CALL, GI_DELAY_READ, GI_DELAY_WRITE = range(5)
(BASIC_ZERO, BASIC_SET, BASIC_COPY, BASIC_ADD, BASIC_ADD_INTO, BASIC_ADD_SCALAR, BASIC_ADD_SCALAR_INTO, BASIC_MULTIPLY,
 BASIC_MULTIPLY_WITH, BASIC_MULTIPLY_SCALAR, BASIC_MULTIPLY_WITH_SCALAR) = range(11)
ARITH_SUB, ARITH_DIV, ARITH_MIN, ARITH_MAX, ARITH_NEG, ARITH_ABS, ARITH_SQRT, ARITH_SIN, ARITH_COS, ARITH_POW = range(10)
SCALAR_NONE, SCALAR_A, SCALAR_B = range(3)
GRAPH_OUT_BUFFERS, GRAPH_OUT_MIX = 0, 1

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




class GraphRef(C.Structure):
    _fields_ = [("kind", C.c_int32), ("index", C.c_uint32)]


class GraphInstr(C.Structure):
    _fields_ = [("op", C.c_int32), ("sub", C.c_int32), ("out", GraphRef), ("in_", GraphRef * 3), ("scalar", C.c_float),
                ("scalar_side", C.c_int32), ("field", C.c_uint32), ("span_table", C.c_uint32)]


class GraphConfig(C.Structure):
    _fields_ = [("voices", C.c_uint32), ("num_temps", C.c_uint32), ("num_outputs", C.c_uint32), ("num_inputs", C.c_uint32),
                ("num_span_tables", C.c_uint32), ("num_fields", C.c_uint32), ("output_mode", C.c_int32), ("channels", C.c_uint32)]


def ptr(a):
    """void* of a numpy array (or None)."""
    if a is None:
        return None
    return a.ctypes.data_as(C.c_void_p)
