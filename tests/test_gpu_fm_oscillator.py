"""Parity of the fmsynth feedback Oscillator (examples/example_fmsynth.zig:26-89) against the oracle.

The phase carried across calls (`t`, wrapped at the end of every call) only sees f32 mul/add/trunc and must be bit-exact.
The output feeds back through `sin`, whose libm bits are unpinned (oracle: glibc sinf, device: its own <= 2 ulp fast path);
with |feedback| * 2 < 1 the recurrence is a contraction and the difference stays at the 1e-6 level (bound asserted: 1e-5
of full scale, the north-star tolerance).  All four waveforms, with and without a phase buffer, random spans with gaps,
two blocks with carried state.
"""
import numpy as np
import pytest

import oracle
from helpers import random_spans
from zang_b200 import _lib as L
from zang_b200 import engine as E

pytestmark = pytest.mark.gpu

BF = 512
TOL = 1e-5


@pytest.mark.parametrize("with_phase", [False, True], ids=["no_phase", "phase_buf"])
def test_fm_oscillator_parity(ctx, with_phase):
    V, NB = 150, 6
    frames = NB * BF
    rng = np.random.default_rng(31 + with_phase)
    params = np.zeros(8, dtype=L.param_dtypes[L.MOD_FM_OSCILLATOR])
    params["sample_rate"] = 48000.0
    params["waveform"] = np.arange(8) % 3  # waveform 3 is discontinuous: test_fm_oscillator_gated_waveform
    params["feedback"] = rng.uniform(0.0, 0.4, 8)
    mod = E.FmOscillator(ctx, V)
    ostate = oracle.state_init(L.MOD_FM_OSCILLATOR, V)
    for block in range(2):
        off, spans = random_spans(rng, V, NB, BF, 8)
        out0 = rng.uniform(-1, 1, size=(frames, V)).astype(np.float32)
        # t grows by freq/sr per frame until the call ends (512 frames): up to ~4 cycles here, so the sine argument stays below
        # ~40 and its own f32 spacing (<= 3.8e-6) below the tolerance; a last-bit difference in the fed-back sample is re-quantised
        # to that spacing, which is the noise floor of the reference's formulation itself
        freq = (rng.uniform(50, 300, size=(1, V)) * (1 + 0.02 * rng.standard_normal((frames, V)))).astype(np.float32)
        phase = rng.uniform(-2, 2, size=(frames, V)).astype(np.float32) if with_phase else None
        want = out0.copy()
        oracle.paint(L.MOD_FM_OSCILLATOR, off, spans, params, ostate, want, freq, phase)
        d_out, d_f = ctx.buf(frames, V, out0), ctx.buf(frames, V, freq)
        d_p = ctx.buf(frames, V, phase) if with_phase else None
        sp = ctx.spans(V, frames, off, spans)
        mod.paint(sp, [d_out], [], params, d_f, d_p)
        got = d_out.download()
        err = float(np.max(np.abs(got.astype(np.float64) - want.astype(np.float64))))
        assert err <= TOL, f"block {block}: max abs error {err}"
        gs = mod.bank.get_state()
        assert np.array_equal(gs["t"].view(np.uint32), ostate["t"].view(np.uint32)), "carried phase differs"
        assert np.max(np.abs(gs["feedback1"] - ostate["feedback1"])) <= TOL and np.max(np.abs(gs["feedback2"] - ostate["feedback2"])) <= TOL
        # keep the two recurrences on the same footing for the next block
        mod.bank.set_state(ostate)
        for b in (d_out, d_f, d_p, sp):
            if b is not None:
                b.free()


def test_fm_oscillator_gated_waveform(ctx):
    """waveform 3 (:76) gates |sin p| by the SIGN of sin(2p): at the zero crossings of sin(2p) where |sin p| = 1 the output
    jumps between 1 and 0, so a last-bit difference between two `sin` implementations flips single samples.  Without feedback
    such flips do not propagate: everything else must agree to 1e-5 and the flips must be rare."""
    V, NB = 128, 4
    frames = NB * BF
    rng = np.random.default_rng(77)
    params = np.array([(48000.0, 3, 0.0)], dtype=L.param_dtypes[L.MOD_FM_OSCILLATOR])
    off = np.arange(V + 1, dtype=np.uint32)
    spans = np.array([(0, frames, 0, 0)] * V, dtype=L.span_dtype)
    freq = np.broadcast_to(rng.uniform(50, 3000, size=(1, V)), (frames, V)).astype(np.float32).copy()
    want = np.zeros((frames, V), dtype=np.float32)
    ostate = oracle.state_init(L.MOD_FM_OSCILLATOR, V)
    oracle.paint(L.MOD_FM_OSCILLATOR, off, spans, params, ostate, want, freq, None)
    mod = E.FmOscillator(ctx, V)
    d_out, d_f = ctx.buf(frames, V), ctx.buf(frames, V, freq)
    mod.paint(ctx.spans(V, frames, off, spans), [d_out], [], params, d_f, None)
    diff = np.abs(d_out.download().astype(np.float64) - want.astype(np.float64))
    flips = int((diff > TOL).sum())
    assert flips <= diff.size // 2000, f"{flips} of {diff.size} samples differ by more than {TOL}"
    assert np.array_equal(mod.bank.get_state()["t"].view(np.uint32), ostate["t"].view(np.uint32))


def test_fm_oscillator_requires_freq(ctx):
    V, frames = 32, BF
    mod = E.FmOscillator(ctx, V)
    off = np.arange(V + 1, dtype=np.uint32)
    spans = np.array([(0, frames, 0, 0)] * V, dtype=L.span_dtype)
    sp = ctx.spans(V, frames, off, spans)
    out = ctx.buf(frames, V)
    with pytest.raises(E.ZangCuError):
        mod.paint(sp, [out], [], np.zeros(1, dtype=L.param_dtypes[L.MOD_FM_OSCILLATOR]), None)
