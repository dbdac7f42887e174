"""The oracle against its committed golden vectors (tests/golden/oracle_golden.json, made by
tests/golden/make_golden.py), plus hand-checkable properties of the restated arithmetic.

Parity of paint() is UNPINNED by the reference (no test, no golden audio, no Zig toolchain here); these tests
keep the oracle stable and check the few values that follow directly from the Zig source.
"""
import json
import os
import sys

import numpy as np

import oracle
from zang_b200 import _lib as L

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_golden  # noqa: E402

GOLD = json.load(open(os.path.join(HERE, "golden", "oracle_golden.json")))


def test_module_outputs_match_golden_hashes():
    got = make_golden.module_cases()
    assert got.keys() == GOLD["modules"].keys()
    bad = [k for k in got if got[k] != GOLD["modules"][k]]
    assert not bad, bad


def test_excerpts_match_golden_values():
    got = make_golden.excerpts()
    for k, v in GOLD["excerpts"].items():
        assert np.array_equal(np.array(got[k], dtype=np.float32).view(np.uint32), np.array(v, dtype=np.float32).view(np.uint32)), k


def test_render_matches_golden_hashes():
    assert make_golden.render_case() == GOLD["render_nice"]


def test_bank_graphs_match_golden_hashes():
    assert make_golden.bank_cases() == GOLD["banks"]


def test_delay_modules_match_golden():
    got = make_golden.delay_cases()
    assert got == GOLD["delay"]
    assert got["simple_delay_3_literal"] == [0.0, 0.0, 0.0, 1.0, 2.0, 3.0, 4.0, 5.0]  # delay.zig: out[i] += in[i - delay_samples]


def one_voice(kind, params, n=64, flags=1, **kw):
    off = np.array([0, 1], dtype=np.uint32)
    spans = np.array([(0, n, flags, 0)], dtype=L.span_dtype)
    out = np.zeros((n, 1), dtype=np.float32)
    st = oracle.state_init(kind, 1, 0)
    oracle.paint(kind, off, spans, np.array([params], dtype=L.param_dtypes[kind]), st, out, **kw)
    return out[:, 0], st


def test_pulse_phase_is_closed_form():
    # PulseOsc.zig:88-89,112: cnt advances by ifreq = u32(2^32/sr * freq) per sample, mod 2^32
    out, st = one_voice(L.MOD_PULSE_OSC, (48000.0, 440.0, 0.5), n=1000)
    ifreq = int(np.float32(np.float32(4294967296.0) / np.float32(48000.0)) * np.float32(440.0))
    assert int(st["cnt"][0]) == (1000 * ifreq) % (1 << 32)
    assert set(np.round(np.abs(out[np.abs(np.abs(out) - 0.7) < 1e-9]), 6)) == {np.float32(0.7)}  # flat parts are +-gain


def test_pulse_out_of_range_frequency_paints_nothing():  # PulseOsc.zig:82-84
    for f in (-1.0, 6000.1):
        out, st = one_voice(L.MOD_PULSE_OSC, (48000.0, f, 0.5))
        assert not out.any() and int(st["cnt"][0]) == 0


def test_envelope_reaches_one_then_decays_to_sustain():
    out, st = one_voice(L.MOD_ENVELOPE, (48000.0, (1, 0.0005), (1, 0.0005), (1, 0.0005), 0.5, 1), n=64)
    # linear attack over 24 samples: t_step = 1/(0.0005*48000) (painter.zig:97)
    step = np.float32(1.0) / (np.float32(0.0005) * np.float32(48000.0))
    t = np.float32(0)
    for i in range(5):
        t = np.float32(t + step)
        assert out[i] == t
    assert out.max() == np.float32(1.0) and out[-1] == np.float32(0.5) and int(st["state"][0]) == L.ENV_SUSTAIN


def test_noise_pink_state_is_not_persisted():  # Noise.zig:68 assigns the wrong way round
    off = np.array([0, 2], dtype=np.uint32)
    spans = np.array([(0, 8, 0, 0), (8, 16, 0, 0)], dtype=L.span_dtype)
    out = np.zeros((16, 1), dtype=np.float32)
    st = oracle.state_init(L.MOD_NOISE, 1, 0)
    oracle.paint(L.MOD_NOISE, off, spans, np.array([(L.NOISE_PINK,)], dtype=L.param_dtypes[L.MOD_NOISE]), st, out)
    # second call restarts the 7-tap filter from zeros: recompute it from the white sequence
    s = np.zeros(4, dtype=np.uint64)
    oracle.orc.orc_xoshiro_seed(0, L.ptr(s))
    fl = np.zeros(16, dtype=np.float32)
    oracle.orc.orc_xoshiro_floats(L.ptr(s), L.ptr(fl), 16)
    white = fl * np.float32(2.0) - np.float32(1.0)
    f32 = np.float32
    for start in (0, 8):
        b = [f32(0)] * 7
        for i in range(start, start + 8):
            w = white[i]
            b[0] = f32(f32(0.99886) * b[0]) + f32(w * f32(0.0555179))
            b[1] = f32(f32(0.99332) * b[1]) + f32(w * f32(0.0750759))
            b[2] = f32(f32(0.96900) * b[2]) + f32(w * f32(0.1538520))
            b[3] = f32(f32(0.86650) * b[3]) + f32(w * f32(0.3104856))
            b[4] = f32(f32(0.55000) * b[4]) + f32(w * f32(0.5329522))
            b[5] = f32(f32(-0.7616) * b[5]) - f32(w * f32(0.0168980))
            acc = b[0]
            for k in range(1, 7):
                acc = f32(acc + b[k])
            acc = f32(acc + f32(w * f32(0.5362)))
            assert out[i, 0] == acc, i
            b[6] = f32(w * f32(0.115926))


def test_sampler_weights_are_swapped():  # Sampler.zig:121-125
    data = np.array([0, 0, 0x00, 0x40, 0x00, 0x20, 0, 0], dtype=np.uint8)  # s16: 0, 16384, 8192, 0 -> 0, .5, .25, 0
    off = np.array([0, 1], dtype=np.uint32)
    spans = np.array([(0, 4, 1, 0)], dtype=L.span_dtype)
    out = np.zeros((4, 1), dtype=np.float32)
    st = oracle.state_init(L.MOD_SAMPLER, 1, 0)
    params = np.array([(96000.0, 0, 0)], dtype=L.param_dtypes[L.MOD_SAMPLER])  # ratio 0.5
    oracle.paint(L.MOD_SAMPLER, off, spans, params, st, out, aux=data, aux_meta=[1, 48000, L.FMT_SIGNED16_LSB])
    # t = 0, .5, 1, 1.5: tfrac = t1 - t = 1, .5, 1, .5 ; s = s0*(1-tfrac) + s1*tfrac
    assert [float(x) for x in out[:, 0]] == [0.5, 0.25, 0.25, 0.375]


def test_mixdown_clamps():  # mixdown.zig:37-45
    mix = np.array([[10.0], [-10.0], [np.nan], [0.5], [-0.5]], dtype=np.float32)
    dst = np.zeros(10, dtype=np.uint8)
    oracle.orc.orc_mix_down(L.ptr(mix), 5, 1, L.AUDIO_SIGNED16_LSB, 1.0, L.ptr(dst))
    assert [int(x) for x in dst.view("<i2")] == [32766, -32767, 0, 16383, -16383]
