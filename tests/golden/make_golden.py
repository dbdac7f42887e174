"""Regenerates tests/golden/oracle_golden.json from the oracle (oracle/zang_oracle.hpp).

The reference pins no paint() output, and it cannot be executed here (Zig, no toolchain), so
these vectors do not pin the oracle to the reference; they pin the oracle to ITSELF across rounds (any edit
that changes a single output bit of any module shows up) and record short literal excerpts a reader can check
by hand against the Zig source.
Run:  python tests/golden/make_golden.py
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import cases  # noqa: E402
import oracle  # noqa: E402
from zang_b200 import _lib as L  # noqa: E402
from zang_b200 import synth  # noqa: E402


def digest(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        if a is not None:
            h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def module_cases():
    out = {}
    for case in cases.CASES:
        name, kind = case[0], case[1]
        rng = cases.case_rng(name)
        st = oracle.state_init(kind, cases.V, 5)
        for blk in range(2):
            off, spans, params, out0, inputs = cases.gen_block(case, rng)
            want = out0.copy()
            oracle.paint(kind, off, spans, params, st, want, *inputs)
            out[f"{name}/block{blk}"] = digest(want, st)
    return out


def excerpts():
    """Short literal vectors: one voice, one full-buffer span, simple params."""
    ex = {}
    off = np.array([0, 1], dtype=np.uint32)
    spans = np.array([(0, 16, 1, 0)], dtype=L.span_dtype)

    def run(kind, params, **kw):
        out = np.zeros((16, 1), dtype=np.float32)
        st = oracle.state_init(kind, 1, 0)
        oracle.paint(kind, off, spans, np.array([params], dtype=L.param_dtypes[kind]), st, out, **kw)
        return [float(x) for x in out[:, 0]]

    ex["pulse 440Hz color 0.3 @48k"] = run(L.MOD_PULSE_OSC, (48000.0, 440.0, 0.3))
    ex["trisaw 440Hz color 0.5 @48k"] = run(L.MOD_TRISAW_OSC, (48000.0, 440.0, 0.5))
    ex["noise white seed 0"] = run(L.MOD_NOISE, (L.NOISE_WHITE,))
    ex["noise pink seed 0"] = run(L.MOD_NOISE, (L.NOISE_PINK,))
    ex["envelope attack cubed 0.0002s"] = run(L.MOD_ENVELOPE, (48000.0, (3, 0.0002), (3, 0.0001), (3, 0.5), 0.8, 1))
    ex["cycle speed 4800"] = run(L.MOD_CYCLE, (48000.0, 4800.0))
    step = np.zeros((16, 1), dtype=np.float32)
    step[4:] = 1.0
    ex["filter lowpass cutoff 0.2 res 0.7 step"] = run(L.MOD_FILTER, (L.FILTER_LOW_PASS, 0.2, 0.7), in0=step)
    return ex


def render_case():
    wl = synth.polyphony_bank(24, 6 * 1024, 48000.0, seed=42, mean_on=0.04, mean_off=0.02)
    mix, vo, _ = oracle.render_nice(wl, 6 * 1024, window=(0, 6 * 1024))
    return {"workload": digest(wl["event_off"], wl["events"], wl["color"], wl["gains"]), "voices": digest(vo), "mix": digest(mix)}


def bank_cases():
    """configs[2..4] graphs (oracle/bank_oracle.hpp) on small seeded workloads.  The FM and sampler chains evaluate libm
    sinf / atanf per sample, so their digests pin the oracle on THIS libm only (values are rounded to 1e-5 before hashing)."""
    def q(a):
        return np.round(a.astype(np.float64) * 1e5).astype(np.int64)
    out = {}
    wl = synth.sfx_bank(20, 4 * 1024, seed=5)
    _, vo, _ = oracle.render_bank(oracle.BANK_SFX, wl, 4 * 1024, 48000.0, first_seed=3)
    out["sfx"] = digest(vo)
    wl = synth.sampler_bank(12, 5 * 1024, 44100.0, seed=11, mean_gap=0.03)
    pcm = synth.pcm_loop(frames=3000, seed=3)
    _, vo, _ = oracle.render_bank(oracle.BANK_SAMPLER, wl, 5 * 1024, 44100.0, pcm=pcm, pcm_meta=(1, 44100, L.FMT_SIGNED16_LSB))
    out["sampler_q1e-5"] = digest(q(vo))
    wl = synth.fm_bank(10, 5 * 1024, 48000.0, seed=21, mean_note=0.04)
    d, i = synth.fm_mod_inputs(10, 5 * 1024)
    _, vo, _ = oracle.render_bank(oracle.BANK_FM, wl, 5 * 1024, 48000.0, in0=d, in1=i)
    out["fm_q1e-4"] = digest(np.round(vo.astype(np.float64) * 1e4).astype(np.int64))
    return out


def delay_cases():
    """src/zang/delay.zig + examples/modules.zig:341-525: hashes on seeded inputs, and the literal SimpleDelay(3) response
    to 1, 2, 3, ... (a pure delay by three painted frames)."""
    from helpers import full_spans, random_spans
    rng = np.random.default_rng(99)
    out = {}
    V, NB, BF = 5, 4, 64
    for kind, name in ((L.MOD_SIMPLE_DELAY, "simple_delay"), (L.MOD_FILTERED_ECHOES, "filtered_echoes")):
        for D in (1, 9, 100):
            off, spans = random_spans(rng, V, NB, BF, 2)
            st = np.zeros(V, dtype=L.state_dtypes[kind])
            line = np.zeros((D, V), dtype=np.float32)
            params = np.array([(0.5, 0.3), (0.8, 0.05)], dtype=L.param_dtypes[L.MOD_FILTERED_ECHOES]) if kind == L.MOD_FILTERED_ECHOES else None
            x = rng.uniform(-1, 1, size=(NB * BF, V)).astype(np.float32)
            o = rng.uniform(-1, 1, size=(NB * BF, V)).astype(np.float32)
            oracle.paint_delay(kind, off, spans, params, st, line, o, x)
            out[f"{name}_D{D}"] = digest(o, line, st)
    x = (np.arange(64, dtype=np.float32) == 0).astype(np.float32).reshape(64, 1)
    left, right = oracle.stereo_echoes(x, 16, 10, 0.5, 0.5)
    out["stereo_echoes_impulse"] = digest(left, right)
    off, spans = full_spans(1, 1, 8)
    st = np.zeros(1, dtype=L.state_dtypes[L.MOD_SIMPLE_DELAY])
    line = np.zeros((3, 1), dtype=np.float32)
    o = np.zeros((8, 1), dtype=np.float32)
    oracle.paint_delay(L.MOD_SIMPLE_DELAY, off, spans, None, st, line, o, np.arange(1, 9, dtype=np.float32).reshape(8, 1))
    out["simple_delay_3_literal"] = [float(v) for v in o[:, 0]]
    # fmsynth Oscillator (examples/example_fmsynth.zig:26-89): libm sinf per sample, hashed after rounding to 1e-5
    V, NB, BF = 4, 3, 64
    off, spans = random_spans(rng, V, NB, BF, 4)
    params = np.array([(48000.0, w, 0.1 * w) for w in range(4)], dtype=L.param_dtypes[L.MOD_FM_OSCILLATOR])
    st = oracle.state_init(L.MOD_FM_OSCILLATOR, V)
    o = np.zeros((NB * BF, V), dtype=np.float32)
    freq = np.full((NB * BF, V), 440.0, dtype=np.float32) * np.arange(1, V + 1, dtype=np.float32)
    oracle.paint(L.MOD_FM_OSCILLATOR, off, spans, params, st, o, freq, None)
    out["fm_oscillator"] = digest(np.round(o.astype(np.float64), 5).astype(np.float32), st["t"])
    return out


def build():
    return {"modules": module_cases(), "excerpts": excerpts(), "render_nice": render_case(), "banks": bank_cases(), "delay": delay_cases()}


if __name__ == "__main__":
    with open(os.path.join(HERE, "oracle_golden.json"), "w") as f:
        json.dump(build(), f, indent=1, sort_keys=True)
    print("written")
