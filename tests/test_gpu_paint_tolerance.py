"""ZCU_PAINT_TOLERANCE: the Filter paint cut into time segments (verified warm-up, else the previous segment's end state).

Bar (BASELINE.json north_star): filter output within 1e-5 of full scale of the reference loop (src/modules/Filter.zig:124-147,
restated in oracle/zang_oracle.hpp).  The warm-up discards the filter's start state, so the comparison is against the
sequential oracle over the WHOLE block, including every segment start.  Cases:
  * damped filters with per-span constants (random notes): segments start from warm-ups;
  * very low cutoff / undamped resonance (res = 1): the warm-up cannot verify, segments chain -> bit-exact;
  * buffer-driven cutoff / resonance;
  * sparse spans (voices silent for whole segments) and voices that never played before a segment (loaded state kept);
  * input aliasing the output: falls back to the exact path.
"""
import numpy as np
import pytest

import oracle
from helpers import assert_bit_exact, full_spans, random_spans, state_equal
from zang_b200 import _lib as L
from zang_b200 import engine as E

pytestmark = pytest.mark.gpu

TOL = 1e-5  # of full scale (inputs are in [-1, 1])
BF = 1024


def run_filter(ctx, V, NB, off, spans, params, x, cutoff=None, res=None, state0=None, mode=L.PAINT_TOLERANCE, alias=False):
    frames = NB * BF
    mod = E.Filter(ctx, V)
    ostate = oracle.state_init(L.MOD_FILTER, V, first_seed=0)
    if state0 is not None:
        ostate[:] = state0
        mod.bank.set_state(ostate)
    out0 = np.zeros((frames, V), dtype=np.float32) if not alias else x.copy()
    want = out0.copy()
    oracle.paint(L.MOD_FILTER, off, spans, params, ostate, want, x if not alias else out0.copy(), cutoff, res)
    d_out = ctx.buf(frames, V, out0)
    d_x = d_out if alias else ctx.buf(frames, V, x)
    d_c = None if cutoff is None else ctx.buf(frames, V, cutoff)
    d_r = None if res is None else ctx.buf(frames, V, res)
    sp = ctx.spans(V, frames, off, spans)
    ctx.set_paint_mode(mode)
    try:
        mod.paint(sp, [d_out], [], params, input=d_x, cutoff=d_c, res=d_r)
    finally:
        ctx.set_paint_mode(L.PAINT_EXACT)
    got = d_out.download()
    gstate = mod.bank.get_state()
    for b in [d_out, sp, d_c, d_r] + ([] if alias else [d_x]):
        if b is not None:
            b.free()
    return got, want, gstate, ostate


def note_params(rng, n, lo=0.02, hi=0.6, types=(L.FILTER_LOW_PASS, L.FILTER_BAND_PASS, L.FILTER_HIGH_PASS, L.FILTER_NOTCH, L.FILTER_ALL_PASS)):
    p = np.zeros(n, dtype=L.param_dtypes[L.MOD_FILTER])
    p["type"] = rng.choice(types, size=n)
    p["cutoff"] = rng.uniform(lo, hi, size=n)
    p["res"] = rng.uniform(0.0, 0.8, size=n)
    return p


def max_err(got, want):
    return float(np.max(np.abs(got.astype(np.float64) - want.astype(np.float64))))


@pytest.mark.parametrize("V,NB", [(256, 64), (1000, 48)])
def test_filter_segments_full_spans(ctx, V, NB):
    rng = np.random.default_rng(V)
    params = note_params(rng, 16)
    off, spans = full_spans(V, NB, BF)
    spans["param"] = rng.integers(0, 16, size=spans.shape[0])
    x = rng.uniform(-1, 1, size=(NB * BF, V)).astype(np.float32)
    s0 = np.zeros(V, dtype=oracle.state_init(L.MOD_FILTER, V, 0).dtype)
    s0["l"] = rng.uniform(-0.5, 0.5, V)
    s0["b"] = rng.uniform(-0.5, 0.5, V)
    got, want, gs, os_ = run_filter(ctx, V, NB, off, spans, params, x, state0=s0)
    assert max_err(got, want) <= TOL
    assert max(np.max(np.abs(gs["l"] - os_["l"])), np.max(np.abs(gs["b"] - os_["b"]))) <= TOL
    # and the exact mode on the same inputs is bit-exact (same kernel, one segment)
    got2, want2, gs2, os2 = run_filter(ctx, V, NB, off, spans, params, x, state0=s0, mode=L.PAINT_EXACT)
    assert_bit_exact(got2, want2, "exact mode")
    assert state_equal(gs2, os2)


def test_filter_segments_random_spans(ctx):
    V, NB = 300, 64
    rng = np.random.default_rng(11)
    params = note_params(rng, 32)
    off, spans = random_spans(rng, V, NB, BF, 32, max_cuts=2, p_empty=0.3, p_gap=0.3)
    x = rng.uniform(-1, 1, size=(NB * BF, V)).astype(np.float32)
    got, want, gs, os_ = run_filter(ctx, V, NB, off, spans, params, x)
    assert max_err(got, want) <= TOL


def test_filter_segments_undamped_chain_is_exact(ctx):
    """res = 1 (no damping) and cutoff 0.001: no window verifies, every segment waits for its predecessor: the result is the
    sequential walk, bit for bit."""
    V, NB = 256, 64
    rng = np.random.default_rng(5)
    params = np.zeros(2, dtype=L.param_dtypes[L.MOD_FILTER])
    params["type"] = [L.FILTER_LOW_PASS, L.FILTER_BAND_PASS]
    params["cutoff"] = [0.001, 0.3]
    params["res"] = [0.2, 1.0]
    off, spans = full_spans(V, NB, BF)
    spans["param"] = np.repeat(rng.integers(0, 2, size=V), NB)  # per voice one of the two rows
    x = rng.uniform(-0.1, 0.1, size=(NB * BF, V)).astype(np.float32)
    got, want, gs, os_ = run_filter(ctx, V, NB, off, spans, params, x)
    assert_bit_exact(got, want, "chained segments")
    assert state_equal(gs, os_)


def test_filter_segments_buffer_params(ctx):
    V, NB = 256, 64
    rng = np.random.default_rng(6)
    params = note_params(rng, 4)
    off, spans = full_spans(V, NB, BF)
    frames = NB * BF
    x = rng.uniform(-1, 1, size=(frames, V)).astype(np.float32)
    t = np.arange(frames, dtype=np.float32)[:, None] / frames
    cutoff = (0.1 + 0.4 * (0.5 + 0.5 * np.sin(40 * t + rng.uniform(0, 6, V)[None, :]))).astype(np.float32)
    res = rng.uniform(0.0, 0.7, size=(frames, V)).astype(np.float32)
    got, want, gs, os_ = run_filter(ctx, V, NB, off, spans, params, x, cutoff=cutoff, res=res)
    assert max_err(got, want) <= TOL


def test_filter_segments_sparse_and_late_voices(ctx):
    """Voices whose first span lies after several segment starts keep their loaded state; voices silent for longer than any
    window get their state through the chain."""
    V, NB = 256, 96
    rng = np.random.default_rng(8)
    params = note_params(rng, 8, lo=0.05)
    rows, off = [], [0]
    for v in range(V):
        n = 0
        first = int(rng.integers(0, NB - 2))
        for b in range(first, NB):
            if v % 3 == 0 and (b - first) % 17 not in (0, 1):
                continue  # long silences
            rows.append((b * BF + (13 if b == first else 0), (b + 1) * BF, 0, int(rng.integers(0, 8))))
            n += 1
        off.append(off[-1] + n)
    spans = np.array(rows, dtype=L.span_dtype)
    off = np.array(off, dtype=np.uint32)
    x = rng.uniform(-1, 1, size=(NB * BF, V)).astype(np.float32)
    s0 = np.zeros(V, dtype=oracle.state_init(L.MOD_FILTER, V, 0).dtype)
    s0["l"] = rng.uniform(-0.5, 0.5, V)
    s0["b"] = rng.uniform(-0.5, 0.5, V)
    got, want, gs, os_ = run_filter(ctx, V, NB, off, spans, params, x, state0=s0)
    assert max_err(got, want) <= TOL


def test_filter_aliasing_falls_back_to_exact(ctx):
    V, NB = 128, 32
    rng = np.random.default_rng(9)
    params = note_params(rng, 4)
    off, spans = full_spans(V, NB, BF)
    x = rng.uniform(-1, 1, size=(NB * BF, V)).astype(np.float32)
    got, want, gs, os_ = run_filter(ctx, V, NB, off, spans, params, x, alias=True)
    assert_bit_exact(got, want, "aliased input")
    assert state_equal(gs, os_)
