"""Odd shapes through the kernels that stage their data in groups: one voice, voice counts around a warp and a CTA, blocks of
one frame, of one frame less and one more than a group, delay lines just long enough for the prefetch queue.

  k_paint_walk          SimpleDelay (sequential form), FilteredEchoes, exact-mode Filter: groups of 32 frames x 32 voices
  SimpleDelayParMod     + k_delay_line_update: chunks of 256 line slots
  k_paint_prefix_ring   TriSawOsc / Cycle with a frequency buffer: groups of 64 frames, forced time segments

Bit-exact against the oracle (outputs, carried state, delay lines); random spans with gaps; voices that never paint.
"""
import numpy as np
import pytest

import oracle
from helpers import assert_bit_exact, random_spans, state_equal
from zang_b200 import _lib as L
from zang_b200 import engine as E

pytestmark = pytest.mark.gpu

SR = 48000.0


def _spans(rng, V, frames, nrows):
    """spans inside a block that is not a whole number of 64-frame buffers: buffers of 64 frames plus a short last one"""
    bf = 64
    nb = max(1, frames // bf)
    if frames < bf:
        off, sp = random_spans(rng, V, 1, frames, nrows, max_cuts=1, p_empty=0.2, p_gap=0.3)
        return off, sp
    off, sp = random_spans(rng, V, nb, bf, nrows, max_cuts=2, p_empty=0.2, p_gap=0.3)
    tail = nb * bf
    if tail == frames:
        return off, sp
    rows, noff = [], [0]
    for v in range(V):  # half of the voices also paint the short last buffer
        mine = [tuple(x) for x in sp[off[v]:off[v + 1]]]
        if rng.random() < 0.5:
            mine.append((tail, frames, int(rng.integers(0, 2)), int(rng.integers(0, nrows))))
        rows += mine
        noff.append(noff[-1] + len(mine))
    spans = np.array(rows, dtype=L.span_dtype) if rows else np.zeros(0, dtype=L.span_dtype)
    return np.array(noff, dtype=np.uint32), spans


@pytest.mark.parametrize("V", [1, 31, 33, 129])
@pytest.mark.parametrize("frames", [1, 63, 65, 1000])
@pytest.mark.parametrize("kind,D,par", [(L.MOD_SIMPLE_DELAY, 320, "0"), (L.MOD_SIMPLE_DELAY, 321, "1"), (L.MOD_SIMPLE_DELAY, 3, "1"),
                                        (L.MOD_FILTERED_ECHOES, 320, "0"), (L.MOD_FILTERED_ECHOES, 500, "0")],
                         ids=["delay-seq-320", "delay-par-321", "delay-par-3", "echoes-320", "echoes-500"])
def test_delay_edge_shapes(ctx, kind, D, par, V, frames, monkeypatch):
    monkeypatch.setenv("ZCU_DELAY_PARALLEL", par)
    rng = np.random.default_rng(V * 131 + frames * 7 + D)
    mod = (E.SimpleDelay if kind == L.MOD_SIMPLE_DELAY else E.FilteredEchoes)(ctx, V, D)
    ostate = np.zeros(V, dtype=L.state_dtypes[kind])
    oline = np.zeros((D, V), dtype=np.float32)
    params = np.zeros(3, dtype=L.param_dtypes[kind])
    if kind == L.MOD_FILTERED_ECHOES:
        params["feedback_volume"] = rng.uniform(0.3, 0.9, 3)
        params["cutoff"] = rng.uniform(0.05, 0.8, 3)
    for block in range(3):  # three calls: index, filter state and line carried; the line wraps for the short delays
        off, spans = _spans(rng, V, frames, 3)
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
        assert state_equal(mod.bank.get_state(), ostate), f"carried state after block {block}"
        assert_bit_exact(mod.bank.get_delay_line(), oline, f"delay line after block {block}")
        for b in (d_out, d_x, sp):
            b.free()


@pytest.mark.parametrize("V", [1, 33, 129])
@pytest.mark.parametrize("frames", [1, 65, 1000])
def test_exact_filter_edge_shapes(ctx, V, frames):
    """exact-mode Filter (one segment) on small banks = the walker kernel; cutoff from a buffer, resonance constant"""
    rng = np.random.default_rng(V * 17 + frames)
    kind = L.MOD_FILTER
    mod = E.Filter(ctx, V)
    ostate = oracle.state_init(kind, V)
    params = np.zeros(4, dtype=L.param_dtypes[kind])
    params["type"] = [L.FILTER_LOW_PASS, L.FILTER_BAND_PASS, L.FILTER_HIGH_PASS, L.FILTER_BYPASS]
    params["cutoff"] = 0.3
    params["res"] = rng.uniform(0.1, 0.8, 4)
    for block in range(2):
        off, spans = _spans(rng, V, frames, 4)
        out0 = rng.uniform(-1, 1, size=(frames, V)).astype(np.float32)
        x = rng.uniform(-1, 1, size=(frames, V)).astype(np.float32)
        cut = rng.uniform(0.01, 0.9, size=(frames, V)).astype(np.float32)
        want = out0.copy()
        oracle.paint(kind, off, spans, params, ostate, want, x, cut, None)
        d_out, d_x, d_c = ctx.buf(frames, V, out0), ctx.buf(frames, V, x), ctx.buf(frames, V, cut)
        sp = ctx.spans(V, frames, off, spans)
        mod.paint(sp, [d_out], [], params, input=d_x, cutoff=d_c, res=None)
        assert_bit_exact(d_out.download(), want, f"block {block}")
        assert state_equal(mod.bank.get_state(), ostate), f"carried state after block {block}"
        for b in (d_out, d_x, d_c, sp):
            b.free()


@pytest.mark.parametrize("V", [1, 33, 130])
@pytest.mark.parametrize("frames,nseg", [(513, 2), (1000, 7), (4096 + 63, 5)])
@pytest.mark.parametrize("kind", [L.MOD_TRISAW_OSC, L.MOD_CYCLE], ids=["trisaw", "cycle"])
def test_phase_walk_edge_shapes(ctx, kind, V, frames, nseg, monkeypatch):
    """frequency / speed buffer: the phase walk over forced segments whose starts are not group aligned, few voices"""
    monkeypatch.setenv("ZCU_PAINT_SEGMENTS", str(nseg))
    rng = np.random.default_rng(V * 29 + frames + nseg)
    mod = E.MODULE_CLASSES[kind](ctx, V)
    ostate = oracle.state_init(kind, V)
    params = np.zeros(3, dtype=L.param_dtypes[kind])
    params["sample_rate"] = [48000.0, 44100.0, 22050.0]
    if kind == L.MOD_TRISAW_OSC:
        params["color"] = [0.1, 0.5, 0.9]
    for block in range(2):
        off, spans = _spans(rng, V, frames, 3)
        out0 = rng.uniform(-1, 1, size=(frames, V)).astype(np.float32)
        drive = rng.uniform(20.0, 3000.0, size=(frames, V)).astype(np.float32)
        want = out0.copy()
        oracle.paint(kind, off, spans, params, ostate, want, drive, None, None)
        d_out, d_f = ctx.buf(frames, V, out0), ctx.buf(frames, V, drive)
        sp = ctx.spans(V, frames, off, spans)
        if kind == L.MOD_TRISAW_OSC:
            mod.paint(sp, [d_out], [], params, freq=d_f)
        else:
            mod.paint(sp, [d_out], [], params, speed=d_f)
        assert_bit_exact(d_out.download(), want, f"block {block}")
        assert state_equal(mod.bank.get_state(), ostate), f"carried state after block {block}"
        for b in (d_out, d_f, sp):
            b.free()


@pytest.mark.parametrize("kind", [L.MOD_PULSE_OSC, L.MOD_DECIMATOR], ids=["pulse-freq-buffer", "decimator"])
def test_second_pass_modules_on_a_large_bank(ctx, kind):
    """Banks with >= 2.5 CTAs per SM from their voices alone: modules whose time segments cost a second pass over an input
    (chained integer scan of PulseOsc's frequency buffer, Decimator's prefix walk) are painted in one segment there
    (zcu_paint.cu: launch).  Two calls, state carried; bit-exact."""
    import cases
    V, frames = 49152 + 96, 600
    rng = np.random.default_rng(kind * 7 + 3)
    mod = E.MODULE_CLASSES[kind](ctx, V)
    ostate = oracle.state_init(kind, V)
    params = cases.make_params(kind, rng, 4, "ctrl")
    for block in range(2):
        # two spans per voice with a gap between them (vectorised: the generic generator loops over voices in Python)
        cuts = np.sort(rng.integers(0, frames - 2, size=(V, 4)), axis=1) + np.arange(4)  # strictly increasing: no empty span
        spans = np.zeros(2 * V, dtype=L.span_dtype)
        spans["start"] = cuts[:, [0, 2]].reshape(-1)
        spans["end"] = cuts[:, [1, 3]].reshape(-1)
        spans["flags"] = rng.integers(0, 2, size=2 * V)
        spans["param"] = rng.integers(0, 4, size=2 * V)
        off = (2 * np.arange(V + 1)).astype(np.uint32)
        out0 = rng.uniform(-1, 1, size=(frames, V)).astype(np.float32)
        lo, hi = (-200.0, 6500.0) if kind == L.MOD_PULSE_OSC else (-1.0, 1.0)
        x = rng.uniform(lo, hi, size=(frames, V)).astype(np.float32)
        want = out0.copy()
        oracle.paint(kind, off, spans, params, ostate, want, x, None, None)
        d_out, d_x = ctx.buf(frames, V, out0), ctx.buf(frames, V, x)
        sp = ctx.spans(V, frames, off, spans)
        if kind == L.MOD_PULSE_OSC:
            mod.paint(sp, [d_out], [], params, freq=d_x)
        else:
            mod.paint(sp, [d_out], [], params, d_x)
        assert_bit_exact(d_out.download(), want, f"block {block}")
        assert state_equal(mod.bank.get_state(), ostate), f"carried state after block {block}"
        for b in (d_out, d_x, sp):
            b.free()
