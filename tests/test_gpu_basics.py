"""zang basics (src/zang/basics.zig:12-78) on [frame][voice] buffers: whole-buffer (vectorised) and
span-restricted forms, bit-exact against the oracle; voice mix and zang.mixDown."""
import numpy as np
import pytest

import oracle
from helpers import assert_bit_exact, assert_close, random_spans
from zang_b200 import _lib as L

pytestmark = pytest.mark.gpu

OPS = ["zero", "set", "copy", "add", "addInto", "addScalar", "addScalarInto", "multiply", "multiplyWith", "multiplyScalar",
       "multiplyWithScalar"]


def call(ctx, op, spans, dest, a, b, s):
    f = getattr(ctx, OPS[op])
    if op == 0:
        f(spans, dest)
    elif op in (1, 6, 10):
        f(spans, dest, s)
    elif op in (2, 4, 8):
        f(spans, dest, a)
    elif op in (3, 7):
        f(spans, dest, a, b)
    else:
        f(spans, dest, a, s)


@pytest.mark.parametrize("shape", [(2048, 64), (1000, 70), (3, 5), (1, 1), (4096, 257)])
@pytest.mark.parametrize("op", range(11), ids=OPS)
def test_basic_full_buffer(ctx, op, shape):
    rng = np.random.default_rng(op * 100 + shape[0])
    frames, V = shape
    d0 = rng.uniform(-2, 2, size=shape).astype(np.float32)
    a = rng.uniform(-2, 2, size=shape).astype(np.float32)
    b = rng.uniform(-2, 2, size=shape).astype(np.float32)
    want = d0.copy()
    oracle.basic(op, None, None, want, a, b, 0.37)
    dd, da, db = ctx.buf(frames, V, d0), ctx.buf(frames, V, a), ctx.buf(frames, V, b)
    call(ctx, op, None, dd, da, db, 0.37)
    assert_bit_exact(dd.download(), want, OPS[op])
    for x in (dd, da, db):
        x.free()


@pytest.mark.parametrize("op", range(11), ids=OPS)
def test_basic_in_spans(ctx, op):
    rng = np.random.default_rng(op + 500)
    V, nb, bf = 70, 3, 1000
    frames = nb * bf
    off, spans = random_spans(rng, V, nb, bf, 1)
    d0 = rng.uniform(-2, 2, size=(frames, V)).astype(np.float32)
    a = rng.uniform(-2, 2, size=(frames, V)).astype(np.float32)
    b = rng.uniform(-2, 2, size=(frames, V)).astype(np.float32)
    want = d0.copy()
    oracle.basic(op, off, spans, want, a, b, -1.25)
    dd, da, db = ctx.buf(frames, V, d0), ctx.buf(frames, V, a), ctx.buf(frames, V, b)
    sp = ctx.spans(V, frames, off, spans)
    call(ctx, op, sp, dd, da, db, -1.25)
    got = dd.download()
    assert_bit_exact(got, want, OPS[op])
    for x in (dd, da, db, sp):
        x.free()


def test_empty_and_ragged(ctx):
    # no spans at all, and voices with no spans next to voices with many
    V, frames = 5, 256
    off = np.zeros(V + 1, dtype=np.uint32)
    sp = ctx.spans(V, frames, off, np.zeros(0, dtype=L.span_dtype))
    d0 = np.ones((frames, V), dtype=np.float32)
    dd = ctx.buf(frames, V, d0)
    ctx.zero(sp, dd)
    assert_bit_exact(dd.download(), d0, "empty span table must not touch the buffer")
    spans = np.array([(0, 256, 0, 0), (0, 1, 0, 0), (1, 2, 0, 0), (255, 256, 0, 0)], dtype=L.span_dtype)
    off = np.array([0, 1, 1, 4, 4, 4], dtype=np.uint32)
    sp2 = ctx.spans(V, frames, off, spans)
    ctx.set(sp2, dd, 7.0)
    want = d0.copy()
    oracle.basic(1, off, spans, want, None, None, 7.0)
    assert_bit_exact(dd.download(), want, "ragged")


@pytest.mark.parametrize("channels", [1, 2])
def test_mix_reduce(ctx, channels):
    rng = np.random.default_rng(9)
    frames, V = 777, 300
    src = rng.uniform(-1, 1, size=(frames, V)).astype(np.float32)
    gains = rng.uniform(0, 0.1, size=(V, channels)).astype(np.float32)
    dst0 = rng.uniform(-1, 1, size=(frames, channels)).astype(np.float32)
    want = dst0.copy()
    oracle.orc.orc_mix_reduce(oracle.ptr(src), frames, V, oracle.ptr(gains), channels, oracle.ptr(want))
    ds, dd = ctx.buf(frames, V, src), ctx.buf(frames, channels, dst0)
    ctx.mix_reduce(ds, gains, channels, dd)
    # the reference adds voices one after another into the shared output; the device tree-sums: the order is
    # not pinned, so the bound is 1e-5 x sum|gain|
    assert_close(dd.download(), want, 1e-5 * float(np.abs(gains).sum(axis=0).max()), "mix")
    again = ctx.buf(frames, channels, dst0)
    ctx.mix_reduce(ds, gains, channels, again)
    assert_bit_exact(again.download(), dd.download(), "mix must be deterministic")


@pytest.mark.parametrize("fmt", [L.AUDIO_SIGNED16_LSB, L.AUDIO_SIGNED8])
def test_mix_down(ctx, fmt):
    rng = np.random.default_rng(11)
    frames, ch = 5000, 2
    mix = rng.uniform(-6, 6, size=(frames, ch)).astype(np.float32)
    mix[5, 0] = np.nan
    mix[6, 1] = np.inf
    mix[7, 0] = -np.inf
    mix[8, 0] = 4.0          # exactly +full scale at vol 0.25 -> clamps to 32766 / 126 (mixdown.zig:37-40)
    mix[9, 0] = -4.0
    want = np.zeros(frames * ch * (2 if fmt == L.AUDIO_SIGNED16_LSB else 1), dtype=np.uint8)
    oracle.orc.orc_mix_down(oracle.ptr(mix), frames, ch, fmt, 0.25, oracle.ptr(want))
    dm = ctx.buf(frames, ch, mix)
    got = ctx.mixDown(dm, fmt, 0.25)
    assert np.array_equal(got, want)
