"""Parity of the delay modules against the oracle restatement of src/zang/delay.zig:7-91 and
examples/modules.zig:341-525.  Bit-exact: the only arithmetic is f32 add/mul and the Filter recurrence.

Cases: delay shorter than a span (the reference serves such a span in several chunks, delay.zig:28-33), delay longer than a
mix buffer, delay_samples = 1, random disjoint spans with gaps (the line only advances on painted frames), two consecutive
blocks (index, filter state and the line itself carried), line checkpoint / resume, StereoEchoes composed from the C ABI.
"""
import numpy as np
import pytest

import oracle
from helpers import assert_bit_exact, full_spans, random_spans, state_equal
from zang_b200 import _lib as L
from zang_b200 import engine as E

pytestmark = pytest.mark.gpu

BF = 512


def _run_blocks(ctx, kind, V, NB, D, span_gen, rng, nrows=4):
    frames = NB * BF
    mod = (E.SimpleDelay if kind == L.MOD_SIMPLE_DELAY else E.FilteredEchoes)(ctx, V, D)
    ostate = np.zeros(V, dtype=L.state_dtypes[kind])
    oline = np.zeros((D, V), dtype=np.float32)
    params = np.zeros(nrows, dtype=L.param_dtypes[kind])
    if kind == L.MOD_FILTERED_ECHOES:
        params["feedback_volume"] = rng.uniform(0.3, 0.9, nrows)
        params["cutoff"] = rng.uniform(0.05, 0.8, nrows)
    for block in range(2):
        off, spans = span_gen()
        out0 = rng.uniform(-1, 1, size=(frames, V)).astype(np.float32)
        x = rng.uniform(-1, 1, size=(frames, V)).astype(np.float32)
        want = out0.copy()
        oracle.paint_delay(kind, off, spans, params, ostate, oline, want, x)
        d_out, d_x = ctx.buf(frames, V, out0), ctx.buf(frames, V, x)
        sp = ctx.spans(V, frames, off, spans)
        if kind == L.MOD_SIMPLE_DELAY:
            mod.paint(sp, [d_out], [], d_x)
        else:
            mod.paint(sp, [d_out], [], params, d_x)
        assert_bit_exact(d_out.download(), want, f"block {block}")
        assert state_equal(mod.bank.get_state(), ostate), f"carried index / filter state differs after block {block}"
        assert_bit_exact(mod.bank.get_delay_line(), oline, f"delay line after block {block}")
        for b in (d_out, d_x, sp):
            b.free()
    return mod


@pytest.mark.parametrize("kind", [L.MOD_SIMPLE_DELAY, L.MOD_FILTERED_ECHOES], ids=["SimpleDelay", "FilteredEchoes"])
@pytest.mark.parametrize("D", [1, 7, 300, 1500])
def test_delay_full_spans(ctx, kind, D):
    V, NB = 133, 6
    rng = np.random.default_rng(D)
    _run_blocks(ctx, kind, V, NB, D, lambda: full_spans(V, NB, BF), rng, nrows=1)


@pytest.mark.parametrize("kind", [L.MOD_SIMPLE_DELAY, L.MOD_FILTERED_ECHOES], ids=["SimpleDelay", "FilteredEchoes"])
def test_delay_random_spans(ctx, kind):
    V, NB, D = 200, 8, 333
    rng = np.random.default_rng(3)
    _run_blocks(ctx, kind, V, NB, D, lambda: random_spans(rng, V, NB, BF, 4), rng)


@pytest.mark.parametrize("kind", [L.MOD_SIMPLE_DELAY, L.MOD_FILTERED_ECHOES], ids=["SimpleDelay", "FilteredEchoes"])
@pytest.mark.parametrize("D,V", [(300, 200), (1500, 133)])
def test_delay_lockstep_line_queue(ctx, kind, D, V, monkeypatch):
    """Large banks run the delay modules in the lockstep kernel, with the line fetched in slot order into a per-thread queue
    behind the tile ring (zcu_paint_kernel.cuh, M::LINE_RING).  Small banks normally take the walker kernel or (SimpleDelay)
    the time-parallel form: both are switched off here so that the lockstep form is the one that runs."""
    monkeypatch.setenv("ZCU_PAINT_NO_WALK", "1")
    monkeypatch.setenv("ZCU_DELAY_PARALLEL", "0")
    rng = np.random.default_rng(D + V)
    NB = 12
    _run_blocks(ctx, kind, V, NB, D, lambda: random_spans(rng, V, NB, BF, 4, max_cuts=2, p_empty=0.1, p_gap=0.3), rng)


@pytest.mark.parametrize("parallel", ["0", "1"], ids=["sequential", "time-parallel"])
@pytest.mark.parametrize("D,NB", [(5, 40), (333, 40), (2500, 40), (2500, 4)], ids=["5", "333", "2500", "2500-longer-than-the-block"])
def test_simple_delay_both_forms(ctx, D, NB, parallel, monkeypatch):
    """SimpleDelay has two device forms with the same results: the sequential walk (every painted frame reads and writes its
    slot) and, for banks too small to fill the GPU that way, time segments (zcu_modules.cuh SimpleDelayParMod: the k-th painted
    frame adds the old line for k < N and the input N painted frames back otherwise; the line is brought up to date by a
    second kernel).  Random spans with gaps, delays shorter than, comparable to and longer than the spans, a long block so
    that the paint is really cut in time, two calls: outputs, index and line bit-equal to the oracle in both forms."""
    monkeypatch.setenv("ZCU_DELAY_PARALLEL", parallel)
    V = 150
    rng = np.random.default_rng(D)
    _run_blocks(ctx, L.MOD_SIMPLE_DELAY, V, NB, D, lambda: random_spans(rng, V, NB, BF, 4, max_cuts=2, p_empty=0.2, p_gap=0.3), rng)


@pytest.mark.parametrize("kind,D,par", [(L.MOD_SIMPLE_DELAY, 400, "0"), (L.MOD_SIMPLE_DELAY, 400, "1"), (L.MOD_SIMPLE_DELAY, 9, "1"),
                                        (L.MOD_FILTERED_ECHOES, 400, "0"), (L.MOD_FILTERED_ECHOES, 9, "0")],
                         ids=["delay-seq", "delay-par", "delay-par-short", "echoes", "echoes-short"])
def test_delay_overwrite_mode(ctx, kind, D, par, monkeypatch):
    """ZCU_OUT_OVERWRITE through every form of the delay modules (walker kernel, time-parallel SimpleDelay, in-place reads for
    short lines): one call = `zang.zero(span, out); module.paint(span, {out}, ...)`; frames outside the spans keep their garbage."""
    monkeypatch.setenv("ZCU_DELAY_PARALLEL", par)
    V, NB = 150, 10
    frames = NB * BF
    rng = np.random.default_rng(900 + D + int(par))
    mod = (E.SimpleDelay if kind == L.MOD_SIMPLE_DELAY else E.FilteredEchoes)(ctx, V, D)
    ostate = np.zeros(V, dtype=L.state_dtypes[kind])
    oline = np.zeros((D, V), dtype=np.float32)
    params = np.zeros(4, dtype=L.param_dtypes[kind])
    if kind == L.MOD_FILTERED_ECHOES:
        params["feedback_volume"] = rng.uniform(0.3, 0.9, 4)
        params["cutoff"] = rng.uniform(0.05, 0.8, 4)
    ctx.set_paint_output(L.OUT_OVERWRITE)
    try:
        for block in range(2):
            off, spans = random_spans(rng, V, NB, BF, 4, max_cuts=2, p_empty=0.15, p_gap=0.3)
            out0 = rng.uniform(-1, 1, size=(frames, V)).astype(np.float32)
            x = rng.uniform(-1, 1, size=(frames, V)).astype(np.float32)
            want = out0.copy()
            oracle.basic(0, off, spans, want)  # zero(span, out)
            oracle.paint_delay(kind, off, spans, params, ostate, oline, want, x)
            d_out, d_x = ctx.buf(frames, V, out0), ctx.buf(frames, V, x)
            sp = ctx.spans(V, frames, off, spans)
            if kind == L.MOD_SIMPLE_DELAY:
                mod.paint(sp, [d_out], [], d_x)
            else:
                mod.paint(sp, [d_out], [], params, d_x)
            assert_bit_exact(d_out.download(), want, f"block {block}")
            assert state_equal(mod.bank.get_state(), ostate)
            assert_bit_exact(mod.bank.get_delay_line(), oline, f"delay line after block {block}")
            for b in (d_out, d_x, sp):
                b.free()
    finally:
        ctx.set_paint_output(L.OUT_ACCUMULATE)


def test_delay_line_checkpoint_resume(ctx):
    V, NB, D = 64, 4, 700
    rng = np.random.default_rng(4)
    mod = _run_blocks(ctx, L.MOD_FILTERED_ECHOES, V, NB, D, lambda: full_spans(V, NB, BF), rng, nrows=1)
    line, st = mod.bank.get_delay_line(), mod.bank.get_state()
    other = E.FilteredEchoes(ctx, V, D)
    other.bank.set_delay_line(line)
    other.bank.set_state(st)
    assert_bit_exact(other.bank.get_delay_line(), line)
    other.bank.reset()  # delay.zig:19-22
    assert not other.bank.get_delay_line().any() and not other.bank.get_state()["index"].any()


def test_stereo_echoes_composite(ctx):
    """examples/example_delay.zig: StereoEchoes(15000) fed buffer by buffer; here 64 independent inputs, shorter delay."""
    V, NB, MAIN = 64, 12, 1501  # odd: HALF_DELAY = 750 by integer division (modules.zig:463)
    frames = NB * BF
    rng = np.random.default_rng(5)
    x = (rng.uniform(-1, 1, size=(frames, V)) * (rng.random((frames, V)) < 0.02)).astype(np.float32)  # sparse clicks
    want_l, want_r = oracle.stereo_echoes(x, BF, MAIN, 0.6, 0.1)
    mod = E.StereoEchoes(ctx, V, MAIN)
    params = np.zeros(1, dtype=mod.Params)
    params["feedback_volume"], params["cutoff"] = 0.6, 0.1
    d_x, d_l, d_r = ctx.buf(frames, V, x), ctx.buf(frames, V), ctx.buf(frames, V)
    temps = [ctx.buf(frames, V), ctx.buf(frames, V)]
    off, spans = full_spans(V, NB, BF)
    sp = ctx.spans(V, frames, off, spans)
    mod.paint(sp, [d_l, d_r], temps, params, d_x)
    assert_bit_exact(d_l.download(), want_l, "left")
    assert_bit_exact(d_r.download(), want_r, "right")
    assert np.abs(want_r).max() > 0.05  # the echoes are really there
