"""Test-side loader of the CPU oracle (oracle/_build/libzang_oracle.so).  TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import subprocess

import numpy as np

from zang_b200 import abi as L  # dtypes / enums only: the oracle never touches libzangcu.so

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ODIR = os.path.join(ROOT, "oracle")
OPATH = os.path.join(ODIR, "_build", "libzang_oracle.so")


def _build():
    subprocess.run(["make", "-C", ODIR], check=True, stdout=subprocess.DEVNULL)


if not os.path.exists(OPATH):
    _build()
orc = C.CDLL(OPATH)

_P, _I, _U, _F, _U64 = C.c_void_p, C.c_int, C.c_uint32, C.c_float, C.c_uint64
orc.orc_basic.argtypes = [_I, _U, _U, _P, _P, _P, _P, _P, _F]
orc.orc_paint.argtypes = [_I, _U, _U, _P, _P, _P, _U, _P, _P, _P, _P, _P, _P, _U64, _P]
orc.orc_state_init.argtypes = [_I, _U, _U, _P]
orc.orc_paint_delay.argtypes = [_I, _U, _U, _P, _P, _P, _U, _P, _P, _U, _P, _P]
orc.orc_stereo_echoes.argtypes = [_U, _U, _U, _U, _F, _F, _P, _P, _P]
orc.orc_cutoff_from_frequency.argtypes = [_F, _F]
orc.orc_cutoff_from_frequency.restype = _F
orc.orc_xoshiro_seed.argtypes = [_U64, _P]
orc.orc_xoshiro_next.argtypes = [_P, _P, _U]
orc.orc_xoshiro_floats.argtypes = [_P, _P, _U]
orc.orc_trigger_run.argtypes = [_P, _U, _U, _P, _P, _U, _U, _P, _U]
orc.orc_trigger_run.restype = _U
orc.orc_poly_dispatch.argtypes = [_U, _P, _P, _P, _P, _U, _P, _P, _P]
orc.orc_note_tracker_consume.argtypes = [_P, _P, _U, _F, _U, _U, _P, _P]
orc.orc_note_tracker_consume.restype = _U
orc.orc_impulse_queue_push.argtypes = [_P, _U, _U, _U]
orc.orc_mix_down.argtypes = [_P, _U, _U, _I, _F, _P]
orc.orc_mix_reduce.argtypes = [_P, _U, _U, _P, _U, _P]
orc.orc_render_nice.argtypes = [_U, _U, _U, _F, _P, _P, _P, _P, _U, _P, _P, _U, _U, _U]
orc.orc_render_nice.restype = C.c_double
orc.orc_hardware_threads.restype = _U

ptr = L.ptr


def state_init(kind, V, first_seed=0):
    dt = L.state_dtypes[kind]
    if dt is None:
        return None
    st = np.zeros(V, dtype=dt)
    assert orc.orc_state_init(kind, V, first_seed, ptr(st)) == 0
    return st


def paint(kind, off, spans, params, state, out, in0=None, in1=None, in2=None, aux=None, aux_meta=None):
    """In-place on `out` ([frames][V] f32) and `state`."""
    frames, V = out.shape
    params = np.ascontiguousarray(params, dtype=L.param_dtypes[kind]).reshape(-1)
    for a in (in0, in1, in2):
        assert a is None or (a.dtype == np.float32 and a.flags.c_contiguous and a.shape == out.shape)
    aux_len = 0
    meta = None
    if kind == L.MOD_CURVE:
        aux = np.ascontiguousarray(aux, dtype=L.curve_node_dtype).reshape(-1)
        aux_len = aux.size
    elif kind == L.MOD_SAMPLER:
        aux = np.ascontiguousarray(aux, dtype=np.uint8)
        aux_len = aux.size
        meta = np.array(aux_meta, dtype=np.int64)
    rc = orc.orc_paint(kind, V, frames, ptr(off), ptr(spans), ptr(params), params.size, ptr(state), ptr(out), ptr(in0), ptr(in1),
                       ptr(in2), ptr(aux), aux_len, ptr(meta))
    assert rc == 0, rc


def paint_delay(kind, off, spans, params, state, line, out, in0):
    """SimpleDelay / FilteredEchoes: in-place on `out`, `state` (index[, l, b] rows) and `line` ([delay_samples][V] f32)."""
    frames, V = out.shape
    assert line.dtype == np.float32 and line.flags.c_contiguous and line.shape[1] == V
    if params is None:
        params = np.zeros(1, dtype=L.param_dtypes[kind])
    params = np.ascontiguousarray(params, dtype=L.param_dtypes[kind]).reshape(-1)
    rc = orc.orc_paint_delay(kind, V, frames, ptr(off), ptr(spans), ptr(params), params.size, ptr(state), ptr(line), line.shape[0], ptr(out),
                             ptr(in0))
    assert rc == 0, rc


def stereo_echoes(inp, buffer_frames, main_delay, feedback_volume, cutoff):
    """StereoEchoes over V mono inputs ([frames][V]), painted buffer by buffer from init(); returns (left, right)."""
    frames, V = inp.shape
    left = np.zeros_like(inp)
    right = np.zeros_like(inp)
    rc = orc.orc_stereo_echoes(V, frames, buffer_frames, main_delay, float(feedback_volume), float(cutoff), ptr(inp), ptr(left), ptr(right))
    assert rc == 0, rc
    return left, right


def basic(op, off, spans, dest, a=None, b=None, scalar=0.0):
    frames, V = dest.shape
    rc = orc.orc_basic(op, V, frames, ptr(off), ptr(spans), ptr(dest), ptr(a), ptr(b), float(scalar))
    assert rc == 0


def render_nice(wl, total_frames, sample_rate=48000.0, buffer_frames=1024, channels=2, want_mix=True, window=None, nthreads=1):
    """Oracle render of a polyphony-bank workload.  Returns (mix or None, voices window or None, seconds)."""
    V = wl["color"].size
    mix = np.zeros((total_frames, channels), dtype=np.float32) if want_mix else None
    vo, first, n = None, 0, 0
    if window is not None:
        first, n = window
        vo = np.zeros((n, V), dtype=np.float32)
    ev = np.ascontiguousarray(wl["events"], dtype=L.note_event_dtype)
    off = np.ascontiguousarray(wl["event_off"], dtype=np.uint32)
    color = np.ascontiguousarray(wl["color"], dtype=np.float32)
    gains = np.ascontiguousarray(wl["gains"], dtype=np.float32)
    secs = orc.orc_render_nice(V, total_frames, buffer_frames, sample_rate, ptr(off), ptr(ev), ptr(color), ptr(gains), channels, ptr(mix),
                               ptr(vo), first, n, nthreads)
    return mix, vo, secs


# ---- the reference's own CPU workload: write_wav of the Toccata song (oracle/song_oracle.hpp) ------------
orc.orc_song_events.argtypes = [C.c_char_p, C.c_size_t, _U, _P, _P, _P, _P, _U]
orc.orc_song_render.argtypes = [C.c_char_p, C.c_size_t, _U64, _F, _P, _P]
orc.orc_song_render.restype = C.c_double
orc.orc_song_render_events.argtypes = [_P, _P, _P, _P, _P, _U64, _F, _P, _P]
orc.orc_song_render_events.restype = C.c_double

song_note_dtype = np.dtype([("t", "<f4"), ("note_id", "<u4"), ("freq", "<f4"), ("note_on", "<i4")])


def song_parse(text):
    """The oracle's doParse: one event array per instrument."""
    if isinstance(text, str):
        text = text.encode("latin-1")
    out = []
    for k in range(3):
        cap = 20000  # example_song.zig:103
        t, ids, fr, on = np.zeros(cap, np.float32), np.zeros(cap, np.uint32), np.zeros(cap, np.float32), np.zeros(cap, np.int32)
        n = orc.orc_song_events(text, len(text), k, ptr(t), ptr(ids), ptr(fr), ptr(on), cap)
        assert n >= 0, "song text does not parse"
        ev = np.zeros(n, dtype=song_note_dtype)
        ev["t"], ev["note_id"], ev["freq"], ev["note_on"] = t[:n], ids[:n], fr[:n], on[:n]
        out.append(ev)
    return out


def song_render(notes, total_frames, sample_rate=48000.0, want_mix=True):
    """write_wav.zig's loop over the parsed events.  Returns (pcm uint8[total*2], mix f32[total] or None, seconds)."""
    counts = np.array([n.size for n in notes], dtype=np.uint32)
    cat = np.concatenate(notes)
    t, ids = np.ascontiguousarray(cat["t"]), np.ascontiguousarray(cat["note_id"])
    fr, on = np.ascontiguousarray(cat["freq"]), np.ascontiguousarray(cat["note_on"])
    pcm = np.zeros(total_frames * 2, dtype=np.uint8)
    mix = np.zeros(total_frames, dtype=np.float32) if want_mix else None
    secs = orc.orc_song_render_events(ptr(counts), ptr(t), ptr(ids), ptr(fr), ptr(on), total_frames, sample_rate, ptr(pcm), ptr(mix))
    assert secs >= 0
    return pcm, mix, secs


# ---- instrument banks of BASELINE.json configs[2..4] (oracle/bank_oracle.hpp) ---------------------------
orc.orc_render_bank.argtypes = [_I, _U, _U, _U, _F, _P, _P, _P, _P, _P, _P, _U64, _U, _U, _I, _P, _P, _U, _P, _U, _P, _P, _U, _U]
orc.orc_render_bank.restype = C.c_double
BANK_SAMPLER, BANK_SFX, BANK_FM = 0, 1, 2


def render_bank(kind, wl, total_frames, sample_rate, buffer_frames=1024, window=None, want_mix=False, pcm=None, pcm_meta=(1, 44100, 1),
                in0=None, in1=None, first_seed=0, channels=2):
    """Oracle render of a configs[2..4] bank.  Returns (mix or None, voices window, seconds)."""
    off = np.ascontiguousarray(wl["event_off"], dtype=np.uint32)
    ev = np.ascontiguousarray(wl["events"], dtype=L.note_event_dtype)
    V = off.size - 1
    first, n = window if window is not None else (0, total_frames)
    vo = np.zeros((n, V), dtype=np.float32)
    mix = np.zeros((total_frames, channels), dtype=np.float32) if want_mix else None
    vf = np.ascontiguousarray(wl["fake_sample_rate"], dtype=np.float32) if "fake_sample_rate" in wl else None
    vi = np.ascontiguousarray(wl["filter_type"], dtype=np.int32) if "filter_type" in wl else None
    nodes = np.ascontiguousarray(wl["nodes"], dtype=L.curve_node_dtype) if "nodes" in wl else None
    gains = np.ascontiguousarray(wl["gains"], dtype=np.float32)
    for a in (in0, in1):
        assert a is None or (a.dtype == np.float32 and a.shape == (total_frames, V) and a.flags.c_contiguous)
    secs = orc.orc_render_bank(kind, V, total_frames, buffer_frames, sample_rate, ptr(off), ptr(ev), ptr(vf), ptr(vi), ptr(nodes), ptr(pcm),
                               0 if pcm is None else pcm.size, pcm_meta[0], pcm_meta[1], pcm_meta[2], ptr(in0), ptr(in1), first_seed,
                               ptr(gains), channels, ptr(mix), ptr(vo), first, n)
    return mix, vo, secs
