"""BASELINE.json configs[2..4] as parity cases: the instrument graphs of zang_b200/banks.py (every reference-style paint
line = one C-ABI module call over all voices) against the oracle's per-voice restatement (oracle/bank_oracle.hpp), on the
same seeded note streams, rendered block by block with module and Trigger state carried across blocks.

Bars: Sampler/Decimator chain and the SFX chain bit-exact up to the per-sample transcendental of Distortion (atanf, 2e-6);
the FM chain is held to a phase-drift bound, see the test.
"""
import numpy as np
import pytest

import oracle
from cases import TRANSCENDENTAL_TOL
from helpers import assert_bit_exact, assert_close
from zang_b200 import _lib as L
from zang_b200 import banks, synth

pytestmark = pytest.mark.gpu


def render_blocks(ctx, bank, total, block, ntemps, extra=None):
    V = bank.voices
    got = np.zeros((total, V), dtype=np.float32)
    for first in range(0, total, block):
        n = min(block, total - first)
        out = ctx.buf(n, V)
        temps = [ctx.buf(n, V) for _ in range(ntemps)]
        ins = []
        if extra is not None:
            ins = [ctx.buf(n, V, np.ascontiguousarray(a[first:first + n])) for a in extra]
        bank.paint_block(first, n, out, temps, *ins)
        got[first:first + n] = out.download()
        for b in [out] + temps + ins:
            b.free()
    return got


@pytest.mark.parametrize("voices,total,block", [(40, 7 * 1024 + 100, 3 * 1024), (33, 4 * 1024, 4 * 1024)])
def test_sampler_bank(ctx, voices, total, block):
    sr = 44100.0
    wl = synth.sampler_bank(voices, total, sr, seed=11, mean_gap=0.03)
    pcm = synth.pcm_loop(frames=3000, seed=3)  # short loop: wrap-around and retriggers inside the test
    _, want, _ = oracle.render_bank(oracle.BANK_SAMPLER, wl, total, sr, pcm=pcm, pcm_meta=(1, 44100, L.FMT_SIGNED16_LSB))
    smp = ctx.sample(pcm, 1, 44100, L.FMT_SIGNED16_LSB)
    bank = banks.SamplerVoiceBank(ctx, wl, smp, sr, total)
    got = render_blocks(ctx, bank, total, block, 2)
    assert np.abs(want).max() > 0.2
    assert_close(got, want, TRANSCENDENTAL_TOL, "sampler bank (atanf per sample)")


def test_sampler_bank_on_the_reference_drumloop(ctx):
    """configs[3] on the reference's own asset (examples/example_sampler.zig:66, examples/drumloop.wav as the fixture
    tests/golden/drumloop.wav): 8 s at 44.1 kHz with rare retriggers, so that some voices play through the whole
    192 437-frame loop and the position check against the BYTE length (Sampler.zig:133-135) happens on the real length."""
    sr = 44100.0
    pcm, name = synth.drumloop_pcm()
    assert name == "examples/drumloop.wav" and pcm.size == 384874
    voices, total, block = 40, 344 * 1024 + 77, 64 * 1024
    wl = synth.sampler_bank(voices, total, sr, seed=19, mean_gap=6.0)
    _, want, _ = oracle.render_bank(oracle.BANK_SAMPLER, wl, total, sr, pcm=pcm, pcm_meta=(1, 44100, L.FMT_SIGNED16_LSB))
    smp = ctx.sample(pcm, 1, 44100, L.FMT_SIGNED16_LSB)
    bank = banks.SamplerVoiceBank(ctx, wl, smp, sr, total)
    got = render_blocks(ctx, bank, total, block, 2)
    assert np.abs(want).max() > 0.2
    assert_close(got, want, TRANSCENDENTAL_TOL, "sampler bank on drumloop.wav (atanf per sample)")
    # the Sampler position of at least one voice passed the end of the loop (192 437 frames) without a retrigger
    st = bank.sampler.bank.get_state()
    assert float(st["t"].max()) > 192437.0, float(st["t"].max())


@pytest.mark.parametrize("jobs,job_frames,block", [(70, 6 * 1024, 2 * 1024), (32, 24576, 8 * 1024)])
def test_sfx_bank(ctx, jobs, job_frames, block):
    sr = 48000.0
    wl = synth.sfx_bank(jobs, job_frames, seed=5)
    _, want, _ = oracle.render_bank(oracle.BANK_SFX, wl, job_frames, sr, first_seed=9)
    bank = banks.SfxBank(ctx, wl, sr, job_frames, first_seed=9)
    got = render_blocks(ctx, bank, job_frames, block, 3)
    assert np.abs(want).max() > 0.05
    assert_bit_exact(got, want, "sfx bank")


FM_TOL = 4e-5


def test_fm_bank(ctx):
    """Portamento, TriSawOsc (buffer freq), Envelope and the products are bit-exact; the two SineOsc evaluate sinf per
    sample (CUDA libm vs glibc, <= 2 ulp each).  The carrier's phase input is the modulator's sinf output times the FM index
    (<= 2), and SineOsc wraps its f32 phase accumulator only at the end of a paint call (SineOsc.zig:40), so inside a 1024-frame
    span `t + phase` reaches ~20 cycles where one f32 ulp is 1.9e-6 cycles = 1.2e-5 rad: a 1-ulp difference in the
    modulator flips that rounding.  Hence the bar here is 2 flips (4e-5 on a voice whose full scale is 2.0: two oscillators
    summed), i.e. 2e-5 of full scale; observed 1.2e-5.  Everything upstream of the sines is checked bit for bit below."""
    sr, voices, total, block = 48000.0, 36, 9 * 1024 + 50, 4 * 1024
    wl = synth.fm_bank(voices, total, sr, seed=21, mean_note=0.04)
    detune, index = synth.fm_mod_inputs(voices, total)
    _, want, _ = oracle.render_bank(oracle.BANK_FM, wl, total, sr, in0=detune, in1=index)
    bank = banks.FmDetunedBank(ctx, wl, sr, total)
    got = render_blocks(ctx, bank, total, block, 6, extra=(detune, index))
    assert np.abs(want).max() > 0.2
    assert_close(got, want, FM_TOL, "fm/detuned bank")
    # the glide really happens: some Portamento spans start from a note that is still held
    bank2 = banks.FmDetunedBank(ctx, wl, sr, total)
    _, _, a_porta, a_env, _ = bank2.span_tables(0, total)
    assert (a_porta["param"] % 2 == 1).sum() > voices and a_env["flags"].sum() > voices
