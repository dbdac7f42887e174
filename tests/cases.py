"""Seeded parity cases shared by the GPU parity tests and the CPU-side oracle golden test."""
import zlib

import numpy as np

from zang_b200 import _lib as L

V = 70
NBUF = 3
SR = 48000.0
TRANSCENDENTAL_TOL = 2e-6  # sinf / atanf: CUDA libm vs glibc, both <= 2 ulp on outputs of magnitude <= 1


def curve_rows(rng, n, kinds=(0, 1, 2, 3)):
    out = np.zeros(n, dtype=L.painter_curve)
    out["kind"] = rng.choice(kinds, size=n)
    out["duration"] = rng.choice([0.001, 0.01, 0.02, 0.1, 0.5], size=n)
    return out


def make_params(kind, rng, n, variant):
    p = np.zeros(n, dtype=L.param_dtypes[kind])
    if kind == L.MOD_SINE_OSC:
        p["sample_rate"] = SR
        p["freq"] = rng.uniform(20, 4000, size=n)
        p["phase"] = rng.uniform(-1, 1, size=n)
    elif kind in (L.MOD_PULSE_OSC, L.MOD_TRISAW_OSC):
        p["sample_rate"] = rng.choice([48000.0, 44100.0], size=n)
        p["freq"] = rng.uniform(20, 5000, size=n)
        p["freq"][rng.random(n) < 0.1] = -5.0       # PulseOsc.zig:82-84: paints nothing
        p["freq"][rng.random(n) < 0.1] = 7000.0     # > sr/8: paints nothing
        p["color"] = rng.choice([0.0, 0.1, 0.3, 0.5, 0.77, 1.0, 1.5, -0.2], size=n)
        if kind == L.MOD_TRISAW_OSC and variant == "const":
            # color 0 gives c1 = gain/0 = inf, only harmless while the phase never enters the (empty) up segment
            # (SURVEY.md A.3); colour >= ~1 makes c2 = -gain/(1-col) huge but finite.  Keep both.
            pass
    elif kind == L.MOD_NOISE:
        p["color"] = rng.integers(0, 2, size=n)
    elif kind == L.MOD_ENVELOPE:
        p["sample_rate"] = SR
        p["attack"] = curve_rows(rng, n)
        p["decay"] = curve_rows(rng, n)
        p["release"] = curve_rows(rng, n)
        p["sustain_volume"] = rng.choice([0.0, 0.5, 0.8, 1.0, 1.2], size=n)
        p["note_on"] = rng.integers(0, 2, size=n)
    elif kind == L.MOD_GATE:
        p["note_on"] = rng.integers(0, 2, size=n)
    elif kind == L.MOD_FILTER:
        p["type"] = rng.integers(0, 6, size=n)
        p["cutoff"] = rng.choice([0.01, 0.05, 0.2, 0.5, 0.8, -0.1], size=n)
        p["res"] = rng.choice([0.0, 0.4, 0.7, 0.95, 1.2, -0.5], size=n)
    elif kind == L.MOD_PORTAMENTO:
        p["sample_rate"] = SR
        p["curve"] = curve_rows(rng, n)
        p["goal"] = rng.uniform(50, 2000, size=n)
        p["note_on"] = rng.integers(0, 2, size=n)
        p["prev_note_on"] = rng.integers(0, 2, size=n)
    elif kind == L.MOD_DECIMATOR:
        p["sample_rate"] = SR
        p["fake_sample_rate"] = rng.choice([6000.0, 1000.0, 47999.0, 48000.0, 96000.0, 0.0, -3.0], size=n)
    elif kind == L.MOD_DISTORTION:
        p["type"] = rng.integers(0, 2, size=n)
        p["ingain"] = rng.uniform(0, 1, size=n)
        p["outgain"] = rng.uniform(0, 1, size=n)
        p["offset"] = rng.uniform(-1, 1, size=n)
    elif kind == L.MOD_CYCLE:
        p["sample_rate"] = SR
        p["speed"] = rng.uniform(0.1, 3000, size=n)
    return p


def buf_input(rng, frames, lo, hi):
    return rng.uniform(lo, hi, size=(frames, V)).astype(np.float32)


# (id, kind, variant, input makers: list of (lo,hi) or None per in0/in1/in2, buffer_frames, exact)
CASES = [
    ("sine cc", L.MOD_SINE_OSC, "cc", [None, None, None], 1024, False),
    ("sine bc", L.MOD_SINE_OSC, "bc", [(20, 4000), None, None], 1024, False),
    ("sine cb", L.MOD_SINE_OSC, "cb", [None, (-1, 1), None], 1000, False),
    ("sine bb", L.MOD_SINE_OSC, "bb", [(20, 4000), (-1, 1), None], 1024, False),
    ("pulse const", L.MOD_PULSE_OSC, "const", [None, None, None], 1024, True),
    ("pulse ctrl", L.MOD_PULSE_OSC, "ctrl", [(-200, 6500), None, None], 1024, True),
    ("trisaw const", L.MOD_TRISAW_OSC, "const", [None, None, None], 1024, True),
    ("trisaw ctrl", L.MOD_TRISAW_OSC, "ctrl", [(20, 3000), None, None], 1000, True),
    ("noise", L.MOD_NOISE, "", [None, None, None], 1024, True),
    ("envelope", L.MOD_ENVELOPE, "", [None, None, None], 1024, True),
    ("envelope short buffers", L.MOD_ENVELOPE, "", [None, None, None], 100, True),
    ("gate", L.MOD_GATE, "", [None, None, None], 1024, True),
    ("filter cc", L.MOD_FILTER, "cc", [(-1, 1), None, None], 1024, True),
    ("filter bc", L.MOD_FILTER, "bc", [(-1, 1), (-0.1, 0.8), None], 1024, True),
    ("filter cb", L.MOD_FILTER, "cb", [(-1, 1), None, (-0.1, 1.1)], 1000, True),
    ("filter bb", L.MOD_FILTER, "bb", [(-1, 1), (-0.1, 0.8), (-0.1, 1.1)], 1024, True),
    ("portamento", L.MOD_PORTAMENTO, "", [None, None, None], 1024, True),
    ("decimator", L.MOD_DECIMATOR, "", [(-1, 1), None, None], 1024, True),
    ("distortion", L.MOD_DISTORTION, "", [(-1, 1), None, None], 1024, False),
    ("cycle const", L.MOD_CYCLE, "c", [None, None, None], 1024, True),
    ("cycle buffer", L.MOD_CYCLE, "b", [(0.1, 3000), None, None], 1000, True),
]




def case_rng(name):
    return np.random.default_rng(zlib.crc32(name.encode()))


def gen_block(case, rng, nrows=9):
    """One block of seeded inputs for a case: (off, spans, params, out0, inputs)."""
    from helpers import random_spans
    name, kind, variant, ins, bf, exact = case
    frames = NBUF * bf
    off, spans = random_spans(rng, V, NBUF, bf, nrows)
    params = make_params(kind, rng, nrows, variant)
    out0 = rng.uniform(-1, 1, size=(frames, V)).astype(np.float32)
    inputs = [None if r is None else buf_input(rng, frames, *r) for r in ins]
    return off, spans, params, out0, inputs
