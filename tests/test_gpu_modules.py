"""Parity of every module paint (C ABI -> CUDA) against the oracle, on the same seeded inputs.

Bar (BASELINE.json north_star): integer / index work and Noise bit-exact; floating point within
1e-5 of full scale.  What is actually asserted is stricter: every module is BIT-EXACT (outputs and
carried state) except where a libm transcendental is evaluated per sample — SineOsc (sinf) and
Distortion/overdrive (atanf) — which are held to 2e-6 (tolerance written at the assert).

Every case runs two consecutive blocks so that state carried across calls (and across span tables)
is covered, with random disjoint spans, gaps, one-frame spans, random note_id_changed flags, outputs
pre-filled with noise (paint accumulates) and a voice count that is not a multiple of the warp size.
"""
import numpy as np
import pytest

import oracle
from helpers import assert_bit_exact, assert_close, random_spans, state_equal
from zang_b200 import _lib as L
from zang_b200 import engine as E

pytestmark = pytest.mark.gpu

from cases import CASES, NBUF, SR, TRANSCENDENTAL_TOL, V, case_rng, gen_block


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_module_overwrite_mode(ctx, case):
    """ZCU_OUT_OVERWRITE: one call = `zang.zero(span, out); module.paint(span, {out}, ...)` (examples/modules.zig:220-246):
    inside the spans the output is what the pair leaves in a buffer of garbage, outside them the garbage is untouched."""
    name, kind, variant, ins, bf, exact = case
    rng = case_rng(name + "/overwrite")
    frames = NBUF * bf
    mod = E.MODULE_CLASSES[kind](ctx, V, first_seed=5)
    ostate = oracle.state_init(kind, V, first_seed=5)
    off, spans, params, out0, inputs = gen_block(case, rng)
    want = out0.copy()
    oracle.basic(0, off, spans, want)  # zero(span, out)
    oracle.paint(kind, off, spans, params, ostate, want, *inputs)
    d_out = ctx.buf(frames, V, out0)
    d_in = [None if a is None else ctx.buf(frames, V, a) for a in inputs]
    sp = ctx.spans(V, frames, off, spans)
    ctx.set_paint_output(L.OUT_OVERWRITE)
    try:
        if kind == L.MOD_SINE_OSC:
            mod.paint(sp, [d_out], [], params, freq=d_in[0], phase=d_in[1])
        elif kind in (L.MOD_PULSE_OSC, L.MOD_TRISAW_OSC):
            mod.paint(sp, [d_out], [], params, freq=d_in[0])
        elif kind == L.MOD_FILTER:
            mod.paint(sp, [d_out], [], params, input=d_in[0], cutoff=d_in[1], res=d_in[2])
        elif kind in (L.MOD_DECIMATOR, L.MOD_DISTORTION):
            mod.paint(sp, [d_out], [], params, input=d_in[0])
        elif kind == L.MOD_CYCLE:
            mod.paint(sp, [d_out], [], params, speed=d_in[0])
        else:
            mod.paint(sp, [d_out], [], params)
    finally:
        ctx.set_paint_output(L.OUT_ACCUMULATE)
    got = d_out.download()
    if exact:
        assert_bit_exact(got, want, f"{name} overwrite")
    else:
        assert_close(got, want, TRANSCENDENTAL_TOL, f"{name} overwrite")
    for b in [d_out, sp] + [x for x in d_in if x is not None]:
        b.free()


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_module_parity(ctx, case):
    name, kind, variant, ins, bf, exact = case
    rng = case_rng(name)
    frames = NBUF * bf
    mod = E.MODULE_CLASSES[kind](ctx, V, first_seed=5)
    ostate = oracle.state_init(kind, V, first_seed=5)
    if ostate is not None:
        assert state_equal(mod.bank.get_state(), ostate), "init() state differs"
    for block in range(2):
        off, spans, params, out0, inputs = gen_block(case, rng)
        want = out0.copy()
        oracle.paint(kind, off, spans, params, ostate, want, *inputs)
        d_out = ctx.buf(frames, V, out0)
        d_in = [None if a is None else ctx.buf(frames, V, a) for a in inputs]
        sp = ctx.spans(V, frames, off, spans)
        if kind == L.MOD_SINE_OSC:
            mod.paint(sp, [d_out], [], params, freq=d_in[0], phase=d_in[1])
        elif kind in (L.MOD_PULSE_OSC, L.MOD_TRISAW_OSC):
            mod.paint(sp, [d_out], [], params, freq=d_in[0])
        elif kind == L.MOD_FILTER:
            mod.paint(sp, [d_out], [], params, input=d_in[0], cutoff=d_in[1], res=d_in[2])
        elif kind in (L.MOD_DECIMATOR, L.MOD_DISTORTION):
            mod.paint(sp, [d_out], [], params, input=d_in[0])
        elif kind == L.MOD_CYCLE:
            mod.paint(sp, [d_out], [], params, speed=d_in[0])
        else:
            mod.paint(sp, [d_out], [], params)
        got = d_out.download()
        if exact:
            assert_bit_exact(got, want, f"{name} block {block}")
        else:
            assert_close(got, want, TRANSCENDENTAL_TOL, f"{name} block {block}")
        if ostate is not None:
            gs = mod.bank.get_state()
            if kind == L.MOD_SINE_OSC and variant in ("cb", "bb", "cc", "bc"):
                assert state_equal(gs, ostate), f"{name}: carried phase differs"  # t only sees +, *, trunc: exact
            else:
                assert state_equal(gs, ostate), f"{name}: carried state differs after block {block}"
        for b in [d_out, sp] + [x for x in d_in if x is not None]:
            b.free()


def test_curve_parity(ctx):
    rng = np.random.default_rng(77)
    bf, frames = 1024, NBUF * 1024
    # three curves of different lengths packed in one nodes array (Curve.zig:29-33: params.curve slice)
    curves = [
        [(1000.0, 0.0), (200.0, 0.1), (100.0, 0.2)],                       # example_laser.zig-style sweep
        [(0.0, 0.0), (1.0, 0.004), (0.5, 0.01), (0.5, 0.03), (0.0, 0.05)],
        [(0.3, 0.0), (0.9, 0.0213), (0.9, 0.0213), (0.1, 0.0426)],          # two nodes on the same frame
    ]
    nodes = np.array([n for c in curves for n in c], dtype=L.curve_node_dtype)
    starts = np.cumsum([0] + [len(c) for c in curves])
    for function in (L.INTERP_LINEAR, L.INTERP_SMOOTHSTEP):
        mod = E.Curve(ctx, V)
        ostate = oracle.state_init(L.MOD_CURVE, V)
        for block in range(3):
            nrows = 3
            params = np.zeros(nrows, dtype=L.param_dtypes[L.MOD_CURVE])
            for i in range(nrows):
                params[i] = (SR, function, starts[i], len(curves[i]))
            off, spans = random_spans(rng, V, NBUF, bf, nrows, p_empty=0.05, p_gap=0.1)
            spans["flags"] = (rng.random(spans.size) < 0.15).astype(np.uint32)
            out0 = rng.uniform(-1, 1, size=(frames, V)).astype(np.float32)
            want = out0.copy()
            oracle.paint(L.MOD_CURVE, off, spans, params, ostate, want, aux=nodes)
            d_out = ctx.buf(frames, V, out0)
            sp = ctx.spans(V, frames, off, spans)
            mod.paint(sp, [d_out], [], params, nodes)
            assert_bit_exact(d_out.download(), want, f"curve fn {function} block {block}")
            assert state_equal(mod.bank.get_state(), ostate)
            d_out.free()
            sp.free()


def pcm_bytes(rng, fmt, frames, channels):
    bps = {L.FMT_UNSIGNED8: 1, L.FMT_SIGNED16_LSB: 2, L.FMT_SIGNED24_LSB: 3, L.FMT_SIGNED32_LSB: 4}[fmt]
    return rng.integers(0, 256, size=frames * channels * bps, dtype=np.uint8)


@pytest.mark.parametrize("fmt", [L.FMT_UNSIGNED8, L.FMT_SIGNED16_LSB, L.FMT_SIGNED24_LSB, L.FMT_SIGNED32_LSB])
def test_sampler_parity(ctx, fmt):
    rng = np.random.default_rng(100 + fmt)
    bf, frames = 1024, NBUF * 1024
    channels, wav_sr = 2, 44100
    data = pcm_bytes(rng, fmt, 5000, channels)
    smp = ctx.sample(data, channels, wav_sr, fmt)
    mod = E.Sampler(ctx, V)
    ostate = oracle.state_init(L.MOD_SAMPLER, V)
    nrows = 10
    for block in range(3):
        params = np.zeros(nrows, dtype=L.param_dtypes[L.MOD_SAMPLER])
        # resample at random speeds (example_sampler.zig:124-125), passthrough (ratio ~ 1), reverse (negative rate, loop only),
        # a channel that does not exist (Sampler.zig:87-89)
        params["sample_rate"] = [48000.0 * 0.5, 48000.0 * 1.5, 44100.0, 44100.0 * 1.00005, -44100.0 * 0.7, -30000.0, 48000.0, 22050.0, 44100.0, 96000.0]
        params["channel"] = [0, 1, 0, 1, 0, 1, 2, 1, 1, 0]
        params["loop"] = [1, 1, 1, 0, 1, 0, 1, 0, 1, 1]
        off, spans = random_spans(rng, V, NBUF, bf, nrows)
        out0 = rng.uniform(-1, 1, size=(frames, V)).astype(np.float32)
        want = out0.copy()
        oracle.paint(L.MOD_SAMPLER, off, spans, params, ostate, want, aux=data, aux_meta=[channels, wav_sr, fmt])
        d_out = ctx.buf(frames, V, out0)
        sp = ctx.spans(V, frames, off, spans)
        mod.paint(sp, [d_out], [], params, smp)
        assert_bit_exact(d_out.download(), want, f"sampler fmt {fmt} block {block}")
        assert state_equal(mod.bank.get_state(), ostate)
        d_out.free()
        sp.free()
    smp.free()


def test_envelope_named_edges(ctx):
    """The edge cases SURVEY.md App. D lists: retrigger mid-attack, release finishing mid-span,
    sustain_volume >= 1 (Envelope.zig:54-58), note-off while idle, instantaneous phases."""
    bf, nb = 256, 8
    frames = bf * nb
    rows = np.zeros(6, dtype=L.param_dtypes[L.MOD_ENVELOPE])
    rows["sample_rate"] = SR
    rows["attack"] = (L.CURVE_CUBED, 0.01)
    rows["decay"] = (L.CURVE_SQUARED, 0.005)
    rows["release"] = (L.CURVE_LINEAR, 0.004)
    rows["sustain_volume"] = [0.8, 0.8, 1.0, 0.8, 0.8, 0.8]
    rows["note_on"] = [1, 0, 1, 1, 0, 1]
    rows[3]["attack"] = (L.CURVE_INSTANTANEOUS, 0.0)
    rows[4]["release"] = (L.CURVE_INSTANTANEOUS, 0.0)
    rows[5]["decay"] = (L.CURVE_INSTANTANEOUS, 0.0)
    # voice 0: note-off while idle, on, retrigger mid-attack (new id), off (release ends mid span), on with sustain 1
    # voice 1: instantaneous attack / decay / release
    v0 = [(0, 100, 0, 1), (100, 256, 1, 0), (256, 300, 0, 0), (300, 512, 1, 0), (512, 768, 0, 1), (768, 1024, 0, 1), (1024, 1280, 1, 2),
          (1280, 1536, 0, 2), (1536, 1792, 0, 1)]
    v1 = [(0, 256, 1, 3), (256, 512, 0, 3), (512, 600, 0, 4), (600, 768, 1, 5), (768, 1024, 0, 5), (1024, 1100, 0, 1)]
    spans = np.array(v0 + v1, dtype=L.span_dtype)
    off = np.array([0, len(v0), len(v0) + len(v1)], dtype=np.uint32)
    want = np.zeros((frames, 2), dtype=np.float32)
    ostate = oracle.state_init(L.MOD_ENVELOPE, 2)
    oracle.paint(L.MOD_ENVELOPE, off, spans, rows, ostate, want)
    assert want[:100, 0].max() == 0.0 and want[100:, 0].max() > 0.5  # the scenario really exercises the machine
    mod = E.Envelope(ctx, 2)
    d_out = ctx.buf(frames, 2)
    sp = ctx.spans(2, frames, off, spans)
    mod.paint(sp, [d_out], [], rows)
    assert_bit_exact(d_out.download(), want, "envelope edges")
    assert state_equal(mod.bank.get_state(), ostate)


def test_envelope_time_segments_ramp_tables(ctx):
    """A long block is painted as time segments whose skip-ahead over the earlier spans reads the Painter's exact ramp clock
    from per-step tables (zcu_modules.cuh RampTab, built on the host by zcu_envelope_ramp_tabs): ramps caught in the middle
    by a segment start, by a continued span (same note, no new curve) and by a retrigger; six distinct ramp steps, so two of
    them have no table and take the binade walk.  Bit-exact against the sequential loop, carried state included."""
    Vn, NB, bf = 150, 64, 1024
    frames = NB * bf
    rng = np.random.default_rng(77)
    rows = np.zeros(8, dtype=L.param_dtypes[L.MOD_ENVELOPE])
    rows["sample_rate"] = SR
    rows["attack"] = (L.CURVE_CUBED, 0.01)
    rows["decay"] = (L.CURVE_CUBED, 0.1)
    rows["release"] = (L.CURVE_CUBED, 0.5)
    rows["sustain_volume"] = 0.8
    rows["note_on"] = [1, 0, 1, 0, 1, 0, 1, 0]
    rows[2]["attack"] = rows[3]["attack"] = (L.CURVE_SQUARED, 0.3)    # a slow attack: segment starts land inside it
    rows[4]["decay"] = rows[5]["decay"] = (L.CURVE_LINEAR, 0.9)
    rows[6]["release"] = rows[7]["release"] = (L.CURVE_LINEAR, 2.0)  # fifth and sixth distinct steps: no table
    rows[6]["attack"] = rows[7]["attack"] = (L.CURVE_LINEAR, 0.07)
    off, spans = random_spans(rng, Vn, NB, bf, 8, max_cuts=2, p_empty=0.15, p_gap=0.3)
    want = np.zeros((frames, Vn), dtype=np.float32)
    ostate = oracle.state_init(L.MOD_ENVELOPE, Vn)
    oracle.paint(L.MOD_ENVELOPE, off, spans, rows, ostate, want)
    assert want.max() > 0.7
    mod = E.Envelope(ctx, Vn)
    d_out = ctx.buf(frames, Vn)
    sp = ctx.spans(Vn, frames, off, spans)
    mod.paint(sp, [d_out], [], rows)
    assert_bit_exact(d_out.download(), want, "envelope in time segments")
    assert state_equal(mod.bank.get_state(), ostate)
    d_out.free()


def test_paint_argument_errors(ctx):
    """Error behaviour of the boundary: int codes, never an abort."""
    from zang_b200._lib import lib, ptr
    import ctypes as C
    out = ctx.buf(64, 4)
    off, spans = np.array([0, 1, 1, 1, 1], dtype=np.uint32), np.array([(0, 64, 0, 3)], dtype=L.span_dtype)
    sp = ctx.spans(4, 64, off, spans)
    p = np.zeros(2, dtype=L.param_dtypes[L.MOD_GATE])
    assert lib.zcu_paint_gate(ctx._h, sp._h, ptr(p), 2, out._h) == L.ERR_INVALID  # span refers to row 3 of 2
    assert b"params row" in lib.zcu_last_error(ctx._h)
    bank = ctx.bank(L.MOD_FILTER, 4)
    pp = np.zeros(4, dtype=L.param_dtypes[L.MOD_PULSE_OSC])
    assert lib.zcu_paint_pulse_osc(ctx._h, bank._h, sp._h, ptr(pp), 4, out._h, None) == L.ERR_INVALID  # wrong bank kind
    other = ctx.buf(32, 4)
    assert lib.zcu_zero(ctx._h, sp._h, other._h) == L.ERR_INVALID  # span table of another shape
    bad = np.array([(10, 5, 0, 0)], dtype=L.span_dtype)
    h = C.c_void_p()
    assert lib.zcu_spans_upload(ctx._h, 4, 64, ptr(off), ptr(bad), C.byref(h)) == L.ERR_INVALID


@pytest.mark.parametrize("color", [L.NOISE_WHITE, L.NOISE_PINK], ids=["white", "pink"])
def test_noise_time_segments_bit_exact(ctx, color):
    """Noise is painted in time segments that start from an exact GF(2) jump of the xoshiro256++ state (zcu_paint.cu:
    ensure_xjump): a long block, random spans with gaps (every voice has its own number of draws before a segment) and spans
    straddling segment starts (pink mode replays the 7-tap filter from the span start) must reproduce the sequential loop."""
    Vn, NB, bf = 70, 96, 1024
    frames = NB * bf
    rng = np.random.default_rng(123 + color)
    off, spans = random_spans(rng, Vn, NB, bf, 1, max_cuts=2, p_empty=0.2, p_gap=0.3)
    params = np.array([(color,)], dtype=L.param_dtypes[L.MOD_NOISE])
    mod = E.Noise(ctx, Vn, first_seed=1000)
    ostate = oracle.state_init(L.MOD_NOISE, Vn, first_seed=1000)
    out0 = rng.uniform(-1, 1, size=(frames, Vn)).astype(np.float32)
    want = out0.copy()
    oracle.paint(L.MOD_NOISE, off, spans, params, ostate, want)
    d_out = ctx.buf(frames, Vn, out0)
    sp = ctx.spans(Vn, frames, off, spans)
    mod.paint(sp, [d_out], [], params)
    assert_bit_exact(d_out.download(), want, "noise in time segments")
    assert state_equal(mod.bank.get_state(), ostate)
    assert ctx.noise_resync_count() == 0
    d_out.free()
    sp.free()


@pytest.mark.parametrize("sr", [48000.0, 44100.0, 96000.0, 22050.0, 8000.0, 192000.0, 44099.99609375])
def test_div_by_constant_is_exact(ctx, sr):
    """freq[i] / sample_rate (TriSawOsc.zig:152) is evaluated without the branch of the IEEE division sequence; the result must
    be the IEEE quotient for every f32 numerator the fast sequence accepts (all 2^32 bit patterns are swept on the device)."""
    import ctypes as C
    from zang_b200._lib import lib
    bad, claimed = C.c_uint64(), C.c_uint64()
    rc = lib.zcu_selftest_div_by_const(ctx._h, C.c_float(sr), C.byref(bad), C.byref(claimed))
    assert rc == 0
    assert bad.value == 0, f"{bad.value} of {claimed.value} quotients differ from x / {sr}"
    assert claimed.value > 3.3e9  # every finite numerator with a quotient in [1e-30, 1e30] takes the fast sequence


def test_pulse_gain_division_is_exact(ctx):
    """gain / utof23(ifreq) (PulseOsc.zig:138) is evaluated per sample without the branch of the IEEE division sequence; it
    must be the IEEE quotient for every denominator the 32-bit phase increment can produce."""
    import ctypes as C
    from zang_b200._lib import lib
    bad, claimed = C.c_uint64(), C.c_uint64()
    assert lib.zcu_selftest_div_numerator(ctx._h, C.c_float(0.7), C.byref(bad), C.byref(claimed)) == 0
    assert bad.value == 0, f"{bad.value} of {claimed.value} quotients differ"
    assert claimed.value == (1 << 23) - 1


RELAY_CASES = [c for c in CASES if c[0] in ("sine bc", "sine bb", "trisaw ctrl", "cycle buffer")]


@pytest.mark.parametrize("Vn", [150, 152], ids=["150v", "152v-16B-copies"])
@pytest.mark.parametrize("nseg", [0, 7, 24], ids=["planned", "7seg", "24seg"])
@pytest.mark.parametrize("case", RELAY_CASES, ids=[c[0] for c in RELAY_CASES])
def test_buffer_frequency_phase_relay(ctx, case, nseg, Vn, monkeypatch):
    """Oscillators driven by a frequency (speed) buffer carry an order-dependent f32 phase sum: no skip-ahead exists, so
    a walk of its own (zcu_paint_kernel.cuh, k_paint_prefix_ring, M::PREFIX_RING) tabulates the phase at every segment start: one warp
    per 32 voices runs the frames state-only through a cp.async ring, then the segments paint in parallel.  A long block (spans that straddle, end on
    and start on segment boundaries; two voice blocks), planned and forced segment counts, two consecutive calls: the carried
    phase is bit-equal to the sequential loop's, TriSawOsc and Cycle outputs bit-exact, SineOsc within the sinf tolerance."""
    name, kind, variant, ins, bf, exact = case
    NB = 24
    frames = NB * bf
    if nseg:
        monkeypatch.setenv("ZCU_PAINT_SEGMENTS", str(nseg))
    rng = np.random.default_rng(4242 + nseg)
    from cases import buf_input, make_params
    mod = E.MODULE_CLASSES[kind](ctx, Vn)
    ostate = oracle.state_init(kind, Vn)
    for block in range(2):
        off, spans = random_spans(rng, Vn, NB, bf, 9, max_cuts=2, p_empty=0.1, p_gap=0.3)
        params = make_params(kind, rng, 9, variant)
        out0 = rng.uniform(-1, 1, size=(frames, Vn)).astype(np.float32)
        inputs = [None if r is None else buf_input(rng, frames, *r)[:, :1].repeat(Vn, axis=1) * rng.uniform(0.5, 1.0, Vn).astype(np.float32) for r in ins]
        inputs = [None if a is None else np.ascontiguousarray(a, dtype=np.float32) for a in inputs]
        want = out0.copy()
        oracle.paint(kind, off, spans, params, ostate, want, *inputs)
        d_out = ctx.buf(frames, Vn, out0)
        d_in = [None if a is None else ctx.buf(frames, Vn, a) for a in inputs]
        sp = ctx.spans(Vn, frames, off, spans)
        if kind == L.MOD_SINE_OSC:
            mod.paint(sp, [d_out], [], params, freq=d_in[0], phase=d_in[1])
        elif kind == L.MOD_TRISAW_OSC:
            mod.paint(sp, [d_out], [], params, freq=d_in[0])
        else:
            mod.paint(sp, [d_out], [], params, speed=d_in[0])
        got = d_out.download()
        if exact:
            assert_bit_exact(got, want, f"{name} relay block {block}")
        else:
            assert_close(got, want, TRANSCENDENTAL_TOL, f"{name} relay block {block}")
        assert state_equal(mod.bank.get_state(), ostate), f"{name}: carried phase differs after block {block}"
        for b in [d_out, sp] + [x for x in d_in if x is not None]:
            b.free()


def _xoshiro_prev(s):
    """state before the last xoshiro256++ step (python ints; the device form is XoshiroDev::prev)"""
    M = (1 << 64) - 1
    rotl = lambda x, k: ((x << k) | (x >> (64 - k))) & M
    s0, s1, s2, s3 = s
    s3a = rotl(s3, 19)
    c = s1 ^ s2
    s1o = (c ^ (c << 17) ^ (c << 34) ^ (c << 51)) & M
    s2a = s2 ^ ((s1o << 17) & M)
    s0o = s0 ^ s3a
    return [s0o, s1o, s2a ^ s0o, s3a ^ s1o]


def _xoshiro_next(s):
    M = (1 << 64) - 1
    rotl = lambda x, k: ((x << k) | (x >> (64 - k))) & M
    s0, s1, s2, s3 = s
    r = (rotl((s0 + s3) & M, 23) + s0) & M
    t = (s1 << 17) & M
    s2 ^= s0
    s3 ^= s1
    s1 ^= s2
    s0 ^= s3
    s2 ^= t
    s3 = rotl(s3, 45)
    return r, [s0, s1, s2, s3]


@pytest.mark.parametrize("color", [L.NOISE_WHITE, L.NOISE_PINK], ids=["white", "pink"])
def test_noise_draws_with_an_empty_upper_word(ctx, color, monkeypatch):
    """Random.float's rare branches (Noise.zig:51 -> std.rand float): a draw whose upper 32 bits are zero (lz in 32..40: no
    refill), a draw below 2^23 (one more draw) and two zero draws in a row (a third draw).  The four-sample form of the
    kernels tests a whole batch for them at once and, when one is there, steps the generator back and takes the batch one
    sample at a time: states are built backwards from such a draw so that it falls on every position of a batch."""
    monkeypatch.setenv("ZCU_PAINT_SEGMENTS", "1")  # the skip-ahead of later segments assumes one draw per sample
    rng = np.random.default_rng(77)
    targets = []
    for variant in range(3):
        for back in range(10):
            s1, s2 = (int(x) for x in rng.integers(1, 1 << 63, size=2, dtype=np.uint64))
            tgt = [[0, s1, s2, 5 + back], [0, s1, s2, 0], [0, 0, s2, 0]][variant]
            r, _ = _xoshiro_next(tgt)
            assert (r >> 32) == 0 and (variant == 0) == (r >= 1 << 23)
            s = tgt
            for _ in range(back):
                s = _xoshiro_prev(s)
            chk = s
            for _ in range(back):
                _, chk = _xoshiro_next(chk)
            assert chk == tgt
            targets.append(s)
    Vn, frames = len(targets) + 3, 96
    mod = E.Noise(ctx, Vn, first_seed=5)
    ostate = oracle.state_init(L.MOD_NOISE, Vn, first_seed=5)
    for v, s in enumerate(targets):
        ostate["s"][v] = np.array(s, dtype=np.uint64)
    mod.bank.set_state(ostate)
    params = np.array([(color,)], dtype=L.param_dtypes[L.MOD_NOISE])
    off = np.arange(Vn + 1, dtype=np.uint32)
    spans = np.array([(v % 3, frames - (v % 5), 0, 0) for v in range(Vn)], dtype=L.span_dtype)
    out0 = rng.uniform(-1, 1, size=(frames, Vn)).astype(np.float32)
    want = out0.copy()
    oracle.paint(L.MOD_NOISE, off, spans, params, ostate, want)
    d_out = ctx.buf(frames, Vn, out0)
    sp = ctx.spans(Vn, frames, off, spans)
    mod.paint(sp, [d_out], [], params)
    assert_bit_exact(d_out.download(), want, "noise with refills")
    assert state_equal(mod.bank.get_state(), ostate)
    d_out.free()
    sp.free()
