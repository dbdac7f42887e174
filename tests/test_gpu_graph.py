"""Fused module graphs (zcu_graph_*): the instrument bodies of BASELINE.json configs[2..4] as ONE generated
kernel each, against (a) the per-module drop-in path on the same banks, bit for bit: outputs and every module's carried
state, block after block; (b) the oracle's per-voice restatement (oracle/bank_oracle.hpp) at the bars of
tests/test_gpu_banks.py; (c) MIX output mode against the voice sum of (a).  Reference for the instruction list:
src/zangscript/codegen.zig:60-121, runtime.zig:445-740; for the bodies: examples/modules.zig, example_sampler.zig:94-118,
example_laser.zig:77-115, example_detuned.zig:61-99."""
import numpy as np
import pytest

import oracle
from cases import TRANSCENDENTAL_TOL
from helpers import assert_bit_exact, assert_close
from zang_b200 import _lib as L
from zang_b200 import banks, engine as E, synth

pytestmark = pytest.mark.gpu


def _states(bank):
    out = {}
    for name in ("noise", "curve", "filter", "env", "sampler", "decimator", "porta", "trisaw", "modulator", "carrier"):
        m = getattr(bank, name, None)
        if m is not None and m.bank is not None:
            out[name] = m.bank.get_state()
    return out


def _render(ctx, bank, total, block, ntemps, fused, extra=None):
    V = bank.voices
    got = np.zeros((total, V), dtype=np.float32)
    for first in range(0, total, block):
        n = min(block, total - first)
        out = ctx.buf(n, V)
        ins = [ctx.buf(n, V, np.ascontiguousarray(a[first:first + n])) for a in (extra or ())]
        if fused:
            bank.paint_block_fused(first, n, out, *ins)
            temps = []
        else:
            temps = [ctx.buf(n, V) for _ in range(ntemps)]
            bank.paint_block(first, n, out, temps, *ins)
        got[first:first + n] = out.download()
        for b in [out] + temps + ins:
            b.free()
    return got


def _same_states(a, b, what):
    assert a.keys() == b.keys()
    for k in a:
        assert a[k].tobytes() == b[k].tobytes(), f"{what}: carried state of {k} differs between the fused and the per-module path"


@pytest.mark.parametrize("jobs,job_frames,block,device_spans", [(70, 6 * 1024, 2 * 1024, False), (33, 24576, 8 * 1024, False),
                                                                (70, 6 * 1024, 2 * 1024, True)])
def test_sfx_graph(ctx, jobs, job_frames, block, device_spans):
    sr = 48000.0
    wl = synth.sfx_bank(jobs, job_frames, seed=5)
    _, want, _ = oracle.render_bank(oracle.BANK_SFX, wl, job_frames, sr, first_seed=9)
    ref = banks.SfxBank(ctx, wl, sr, job_frames, first_seed=9)
    unfused = _render(ctx, ref, job_frames, block, 3, fused=False)
    bank = banks.SfxBank(ctx, wl, sr, job_frames, first_seed=9)
    if device_spans:
        bank.use_device_spans(wl)
    g = bank.build_graph()
    assert "NoiseMod m0" in g.source and "FilterMod<true, false>" in g.source
    l0 = ctx.launch_count
    got = _render(ctx, bank, job_frames, block, 3, fused=True)
    assert ctx.launch_count - l0 == (-(-job_frames // block)) * (4 if device_spans else 1), "one kernel launch per block (+3 for the device span table)"
    assert np.abs(want).max() > 0.05
    assert_bit_exact(got, unfused, "sfx graph vs per-module path")
    assert_bit_exact(got, want, "sfx graph vs oracle")
    _same_states(_states(bank), _states(ref), "sfx")


@pytest.mark.parametrize("voices,total,block,device_spans,segments", [(40, 7 * 1024 + 100, 3 * 1024, False, 1), (33, 4 * 1024, 4 * 1024, False, 0),
                                                                      (40, 7 * 1024 + 100, 3 * 1024, True, 5), (70, 16 * 1024, 16 * 1024, False, 0)])
def test_sampler_graph(ctx, voices, total, block, device_spans, segments):
    """segments: time segments of the fused kernel (0 = sized to the GPU, 1 = off, n = forced): Sampler skips ahead in closed
    form, the Decimator by its counter plus a silent replay window that re-latches its held value.  Same bits either way."""
    sr = 44100.0
    wl = synth.sampler_bank(voices, total, sr, seed=11, mean_gap=0.03)
    pcm = synth.pcm_loop(frames=3000, seed=3)
    _, want, _ = oracle.render_bank(oracle.BANK_SAMPLER, wl, total, sr, pcm=pcm, pcm_meta=(1, 44100, L.FMT_SIGNED16_LSB))
    smp = ctx.sample(pcm, 1, 44100, L.FMT_SIGNED16_LSB)
    ref = banks.SamplerVoiceBank(ctx, wl, smp, sr, total)
    unfused = _render(ctx, ref, total, block, 2, fused=False)
    bank = banks.SamplerVoiceBank(ctx, wl, smp, sr, total)
    if device_spans:
        bank.use_device_spans(wl)
    g = bank.build_graph()
    g.set_segments(segments)
    got = _render(ctx, bank, total, block, 2, fused=True)
    info = g.segment_info()
    assert info["failed"] == 0, info
    if segments != 1:
        assert info["segments"] > 1 and info["replay_frames"] >= 44100 // 6000 + 2, info
    assert np.abs(want).max() > 0.2
    assert_bit_exact(got, unfused, "sampler graph vs per-module path")
    assert_close(got, want, TRANSCENDENTAL_TOL, "sampler graph vs oracle (atanf per sample)")
    _same_states(_states(bank), _states(ref), "sampler")


@pytest.mark.parametrize("device_spans", [False, True])
def test_fm_graph(ctx, device_spans):
    sr, voices, total, block = 48000.0, 36, 9 * 1024 + 50, 4 * 1024
    wl = synth.fm_bank(voices, total, sr, seed=21, mean_note=0.04)
    detune, index = synth.fm_mod_inputs(voices, total)
    _, want, _ = oracle.render_bank(oracle.BANK_FM, wl, total, sr, in0=detune, in1=index)
    ref = banks.FmDetunedBank(ctx, wl, sr, total)
    unfused = _render(ctx, ref, total, block, 6, fused=False, extra=(detune, index))
    bank = banks.FmDetunedBank(ctx, wl, sr, total)
    if device_spans:  # Trigger replay and the prev_note_on views on the GPU (zcu_event_bank_*)
        bank.use_device_spans(wl)
    g = bank.build_graph()
    assert "SineMod<true, true>" in g.source and "TriSawMod<true>" in g.source
    got = _render(ctx, bank, total, block, 6, fused=True, extra=(detune, index))
    assert np.abs(want).max() > 0.2
    assert_bit_exact(got, unfused, "fm graph vs per-module path")
    assert_close(got, want, 4e-5, "fm graph vs oracle (bar of tests/test_gpu_banks.py::test_fm_bank)")
    _same_states(_states(bank), _states(ref), "fm")


def test_graph_mix_mode_equals_the_voice_sum(ctx):
    """MIX output: output 0 of every job x pan gains into a stereo mix, against the per-module path + zcu_mix_ordered-style
    host sum (tolerance 1e-5 x sum|gain|: the voice summation order differs)."""
    sr, jobs, frames = 48000.0, 200, 5 * 1024 + 17
    wl = synth.sfx_bank(jobs, frames, seed=8)
    ref = banks.SfxBank(ctx, wl, sr, frames, first_seed=3)
    voices_out = _render(ctx, ref, frames, frames, 3, fused=False)
    want = voices_out.astype(np.float64) @ wl["gains"].astype(np.float64)
    bank = banks.SfxBank(ctx, wl, sr, frames, first_seed=3)
    bank.build_graph(output_mode=L.GRAPH_OUT_MIX, channels=2, gains=wl["gains"])
    mix = ctx.buf(frames, 2)
    ctx.zero(None, mix)
    bank.paint_block_fused(0, frames, mix)
    got = mix.download()
    tol = 1e-5 * float(np.abs(wl["gains"]).sum(axis=0).max())
    assert float(np.abs(want).max()) > 0.05
    assert float(np.abs(got - want).max()) <= tol, float(np.abs(got - want).max())
    _same_states(_states(bank), _states(ref), "sfx mix")
    # accumulates like every zang paint: a second block adds to what is there
    bank2 = banks.SfxBank(ctx, wl, sr, frames, first_seed=3)
    bank2.build_graph(output_mode=L.GRAPH_OUT_MIX, channels=2, gains=wl["gains"])
    bank2.paint_block_fused(0, frames, mix)
    assert float(np.abs(mix.download() - 2 * want).max()) <= 2 * tol


def test_graph_arith_and_span_gaps(ctx):
    """The script runtime's overwriting arithmetic (runtime.zig:528-700) and basics over spans with gaps: frames outside the
    spans keep the output's previous contents; inside, out = max(|a - 0.25|, b / 2) + 0.5 built from five instructions."""
    V, frames = 70, 3000
    rng = np.random.default_rng(4)
    a = rng.standard_normal((frames, V)).astype(np.float32)
    bb = rng.standard_normal((frames, V)).astype(np.float32)
    off, spans = [0], []
    for v in range(V):
        pos = int(rng.integers(0, 50))
        while pos < frames - 10:
            ln = int(rng.integers(1, 400))
            end = min(frames, pos + ln, (pos // 1024 + 1) * 1024)
            spans.append((pos, end, 0, 0))
            pos = end + int(rng.integers(0, 3)) * int(rng.integers(0, 200))
        off.append(len(spans))
    spans = np.array(spans, dtype=L.span_dtype)
    off = np.array(off, dtype=np.uint32)
    b = E.GraphBuilder()
    T, OUT, IN = E.T, E.OUT, E.IN
    b.arith(L.ARITH_SUB, T(0), IN(0), 0.25)
    b.arith(L.ARITH_ABS, T(0), T(0))
    b.arith(L.ARITH_DIV, T(1), IN(1), 2.0)
    b.arith(L.ARITH_MAX, T(0), T(0), T(1))
    b.addScalar(OUT(0), T(0), 0.5)
    cfg = E.graph_config(V, b, 2, num_inputs=2)
    g = E.Graph(ctx, cfg, b)
    prior = rng.standard_normal((frames, V)).astype(np.float32)
    out = ctx.buf(frames, V, prior)
    sp = ctx.spans(V, frames, off, spans)
    g.paint([sp], [out], [ctx.buf(frames, V, a), ctx.buf(frames, V, bb)])
    got = out.download()
    want = prior.copy()
    for v in range(V):
        for k in range(off[v], off[v + 1]):
            s, e = int(spans[k]["start"]), int(spans[k]["end"])
            t0 = np.abs(a[s:e, v] - np.float32(0.25))
            t1 = bb[s:e, v] / np.float32(2.0)
            m = np.where(t0 > t1, t0, t1)
            want[s:e, v] = want[s:e, v] + (m + np.float32(0.5))
    assert_bit_exact(got, want, "arith graph")


def test_graph_time_segments_of_a_skippable_instrument(ctx):
    """PulseOsc (constant frequency) x Envelope with Trigger spans: every module has an exact skip-ahead, so the fused kernel is
    cut in time.  Forced to 7 segments, sized to the GPU and uncut: identical bits, identical carried state."""
    sr, V, total = 48000.0, 90, 11 * 1024 + 333
    wl = synth.polyphony_bank(V, total, sr, seed=3, mean_on=0.05, mean_off=0.02)
    from zang_b200.events import BankTrigger
    off, spans = BankTrigger(wl["event_off"], wl["events"], total, 1024).spans(0, total)
    ev = wl["events"]
    pp = np.zeros(ev.size, dtype=L.param_dtypes[L.MOD_PULSE_OSC])
    pp["sample_rate"], pp["freq"], pp["color"] = sr, ev["freq"], 0.3
    ep = np.zeros(ev.size, dtype=L.param_dtypes[L.MOD_ENVELOPE])
    ep["sample_rate"], ep["attack"], ep["decay"], ep["release"] = sr, (L.CURVE_CUBED, 0.01), (L.CURVE_SQUARED, 0.05), (L.CURVE_LINEAR, 0.1)
    ep["sustain_volume"], ep["note_on"] = 0.7, ev["note_on"]
    results = []
    for segments in (1, 7, 0):
        b = E.GraphBuilder()
        b.zero(E.T(0))
        f_osc = b.call(L.MOD_PULSE_OSC, E.T(0))
        b.zero(E.T(1))
        f_env = b.call(L.MOD_ENVELOPE, E.T(1))
        b.multiply(E.OUT(0), E.T(0), E.T(1))
        b.zero(E.T(0))
        f_cur = b.call(L.MOD_CURVE, E.T(0))  # skip-ahead: begin_span per earlier call, exact running x / y inside the straddling one
        b.addInto(E.OUT(0), E.T(0))
        b.zero(E.T(1))
        f_noi = b.call(L.MOD_NOISE, E.T(1))  # skip-ahead: the exact GF(2) jump of xoshiro256, pink spans replayed from their start
        b.multiplyScalar(E.OUT(0), E.T(1), 0.125)
        g = E.Graph(ctx, E.graph_config(V, b, 2), b)
        osc, env, cur, noi = E.PulseOsc(ctx, V), E.Envelope(ctx, V), E.Curve(ctx, V), E.Noise(ctx, V, first_seed=11)
        npar = np.zeros(ev.size, dtype=L.param_dtypes[L.MOD_NOISE])
        npar["color"] = (np.arange(ev.size) // 2) % 2  # white and pink rows
        g.set_field(f_noi, noi.bank, npar)
        g.set_field(f_osc, osc.bank, pp)
        g.set_field(f_env, env.bank, ep)
        cp = np.zeros(ev.size, dtype=L.param_dtypes[L.MOD_CURVE])
        cp["sample_rate"], cp["node_count"] = sr, 4
        cp["function"] = np.arange(ev.size) % 2  # linear and smoothstep rows
        g.set_field(f_cur, cur.bank, cp)
        g.set_curve_nodes(f_cur, np.array([(0.0, 0.0), (0.5, 0.013), (-0.3, 0.031), (0.2, 0.09)], dtype=L.curve_node_dtype))
        g.set_segments(segments)
        out = ctx.buf(total, V)
        ctx.zero(None, out)
        sp = ctx.spans(V, total, off, spans)
        g.paint([sp], [out])
        info = g.segment_info()
        assert info["failed"] == 0
        assert (info["segments"] == 1) == (segments == 1), info
        results.append((out.download(), osc.bank.get_state().tobytes(), env.bank.get_state().tobytes() + cur.bank.get_state().tobytes() + noi.bank.get_state().tobytes()))
        g.free()
    assert float(np.abs(results[0][0]).max()) > 0.1
    for r in results[1:]:
        assert_bit_exact(r[0], results[0][0], "segmented graph vs one segment")
        assert r[1] == results[0][1] and r[2] == results[0][2], "carried state differs"


def test_graph_rejects_bad_lists(ctx):
    b = E.GraphBuilder()
    b.call(L.MOD_FILTER, E.T(0))  # no input buffer
    with pytest.raises(E.ZangCuError):
        E.Graph(ctx, E.graph_config(8, b, 1), b)
    b = E.GraphBuilder()
    b.zero(E.T(3))  # temp out of range
    with pytest.raises(E.ZangCuError):
        E.Graph(ctx, E.graph_config(8, b, 1), b)


@pytest.mark.parametrize("D", [1, 7, 300])
def test_graph_delay_instructions_equal_simple_delay_and_the_oracle(ctx, D):
    """The script runtime's `delay` block (runtime.zig:704-721) as DELAY_READ / DELAY_WRITE: with the block's body
    `out += feedback_temp; feedback_out = input` it is SimpleDelay.paint (examples/modules.zig:362-386).  Random spans with gaps
    (the line only advances on painted frames), two blocks with the line and its index carried: output, index and line
    bit-exact against the oracle restatement of src/zang/delay.zig."""
    from helpers import random_spans, state_equal
    V, frames = 70, 3 * 512
    rng = np.random.default_rng(D)
    b = E.GraphBuilder()
    fd = b.delay_field()
    b.delay_read(fd, E.T(0))
    b.addInto(E.OUT(0), E.T(0))
    b.delay_write(fd, E.IN(0))
    g = E.Graph(ctx, E.graph_config(V, b, 1, num_inputs=1), b)
    bank = E.DelayBank(ctx, L.MOD_SIMPLE_DELAY, V, D)
    g.set_field(fd, bank, None)
    ostate = np.zeros(V, dtype=L.state_dtypes[L.MOD_SIMPLE_DELAY])
    oline = np.zeros((D, V), dtype=np.float32)
    for block in range(2):
        off, spans = random_spans(rng, V, 3, 512, 1)
        out0 = rng.uniform(-1, 1, size=(frames, V)).astype(np.float32)
        x = rng.uniform(-1, 1, size=(frames, V)).astype(np.float32)
        want = out0.copy()
        oracle.paint_delay(L.MOD_SIMPLE_DELAY, off, spans, None, ostate, oline, want, x)
        d_out, d_x = ctx.buf(frames, V, out0), ctx.buf(frames, V, x)
        sp = ctx.spans(V, frames, off, spans)
        g.paint([sp], [d_out], [d_x])
        assert_bit_exact(d_out.download(), want, f"delay graph block {block}")
        assert state_equal(bank.get_state(), ostate), "carried delay index differs"
        assert_bit_exact(bank.get_delay_line(), oline, f"delay line after block {block}")
    g.free()


def test_graph_feedback_echo_and_libm_arith(ctx):
    """A feedback echo written as a script would (delay block whose body feeds its own output back): y = x + 0.6 * delayed(y),
    against a plain f32 loop; then sin / cos / pow of the script runtime's arithmetic within the libm tolerance."""
    V, frames, D = 40, 2048, 37
    rng = np.random.default_rng(12)
    x = rng.uniform(-1, 1, size=(frames, V)).astype(np.float32)
    b = E.GraphBuilder()
    T, OUT, IN = E.T, E.OUT, E.IN
    fd = b.delay_field()
    b.delay_read(fd, T(0))                 # feedback_temp = delayed y
    b.zero(T(1))
    b.multiplyScalar(T(1), T(0), 0.6)      # feedback_out = 0 + delayed * 0.6
    b.addInto(T(1), IN(0))                 #              + x
    b.delay_write(fd, T(1))
    b.copy(OUT(0), T(1))
    b.arith(L.ARITH_SIN, T(0), T(1))
    b.arith(L.ARITH_COS, T(1), T(1))
    b.arith(L.ARITH_POW, T(1), 2.0, T(1))  # 2 ^ cos(y)
    b.copy(OUT(1), T(0))
    b.copy(OUT(2), T(1))
    g = E.Graph(ctx, E.graph_config(V, b, 2, num_outputs=3, num_inputs=1), b)
    bank = E.DelayBank(ctx, L.MOD_SIMPLE_DELAY, V, D)
    g.set_field(fd, bank, None)
    sp = ctx.spans_full(V, frames // 1024, 1024)
    outs = [ctx.buf(frames, V) for _ in range(3)]
    g.paint([sp], outs, [ctx.buf(frames, V, x)])
    y = np.zeros((frames, V), dtype=np.float32)
    line = np.zeros((D, V), dtype=np.float32)
    for f in range(frames):
        d = np.float32(0.0) + line[f % D]
        y[f] = (np.float32(0.0) + d * np.float32(0.6)) + x[f]
        line[f % D] = y[f]
    assert_bit_exact(outs[0].download(), y, "feedback echo")
    assert_close(outs[1].download(), np.sin(y.astype(np.float64)).astype(np.float32), 2e-6, "sin")
    assert_close(outs[2].download(), np.power(2.0, np.cos(y.astype(np.float64))).astype(np.float32), 4e-6, "pow(2, cos)")
    g.free()


def test_graph_filter_with_constant_coefficients_in_time_segments(ctx):
    """The shape of examples/modules.zig's NiceInstrument as a graph: PulseOsc -> Filter (cutoff constant per note) -> x Envelope.
    In tolerance mode the graph is cut into time segments sized to the GPU; the warm-up window comes from the Filter rows
    (the frames after which the homogeneous map of the slowest row is below 1e-8), so no window has to be lengthened.
    1e-5 of full scale against one segment (exact mode); PulseOsc and Envelope states bit-equal."""
    sr, V, total = 48000.0, 90, 40 * 1024
    wl = synth.polyphony_bank(V, total, sr, seed=3, mean_on=0.08, mean_off=0.03)
    from zang_b200.events import BankTrigger
    off, spans = BankTrigger(wl["event_off"], wl["events"], total, 1024).spans(0, total)
    ev = wl["events"]
    pp = np.zeros(ev.size, dtype=L.param_dtypes[L.MOD_PULSE_OSC])
    pp["sample_rate"], pp["freq"], pp["color"] = sr, ev["freq"], 0.3
    fp = np.zeros(ev.size, dtype=L.param_dtypes[L.MOD_FILTER])
    fp["type"] = L.FILTER_LOW_PASS
    fp["cutoff"] = [E.Filter.cutoffFromFrequency(float(f) * 8.0, sr) for f in ev["freq"]]  # modules.zig:236
    fp["res"] = 0.7
    ep = np.zeros(ev.size, dtype=L.param_dtypes[L.MOD_ENVELOPE])
    ep["sample_rate"], ep["attack"], ep["decay"], ep["release"] = sr, (L.CURVE_CUBED, 0.01), (L.CURVE_CUBED, 0.1), (L.CURVE_CUBED, 0.3)
    ep["sustain_volume"], ep["note_on"] = 0.8, ev["note_on"]
    results = []
    for mode in (L.PAINT_EXACT, L.PAINT_TOLERANCE):
        b = E.GraphBuilder()
        b.zero(E.T(0))
        f_osc = b.call(L.MOD_PULSE_OSC, E.T(0))
        b.multiplyWithScalar(E.T(0), 2.5)
        b.zero(E.T(1))
        f_flt = b.call(L.MOD_FILTER, E.T(1), (E.T(0),))
        b.zero(E.T(2))
        f_env = b.call(L.MOD_ENVELOPE, E.T(2))
        b.multiply(E.OUT(0), E.T(1), E.T(2))
        g = E.Graph(ctx, E.graph_config(V, b, 3), b)
        osc, flt, env = E.PulseOsc(ctx, V), E.Filter(ctx, V), E.Envelope(ctx, V)
        g.set_field(f_osc, osc.bank, pp)
        g.set_field(f_flt, flt.bank, fp)
        g.set_field(f_env, env.bank, ep)
        out = ctx.buf(total, V)
        ctx.zero(None, out)
        sp = ctx.spans(V, total, off, spans)
        ctx.set_paint_mode(mode)
        try:
            g.paint([sp], [out])
            info = g.segment_info()
            retries = g.warm_retries()
        finally:
            ctx.set_paint_mode(L.PAINT_EXACT)
        assert info["failed"] == 0 and retries == 0, (info, retries)
        assert (info["segments"] > 1) == (mode == L.PAINT_TOLERANCE), info
        if mode == L.PAINT_TOLERANCE:
            assert 8 < info["replay_frames"] <= 16384, info
        results.append((out.download(), osc.bank.get_state().tobytes() + env.bank.get_state().tobytes(), flt.bank.get_state().view(np.float32).copy()))
        g.free()
    assert float(np.abs(results[0][0]).max()) > 0.2
    assert_close(results[1][0], results[0][0], 1e-5, "Filter graph in time segments vs one segment")
    assert results[1][1] == results[0][1], "carried PulseOsc / Envelope state"
    assert np.abs(results[1][2] - results[0][2]).max() <= 1e-5, "carried Filter state"


def test_graph_decimator_replay_and_filter_warmup_together(ctx):
    """PulseOsc -> Decimator (held value re-latched in a replay window) -> Filter (constant coefficients: warm-up window from its
    rows) -> x Envelope: the silent window before a segment serves both, and is the longer of the two.  Tolerance mode, forced
    and GPU-sized segments, against one segment in exact mode: 1e-5; PulseOsc / Decimator / Envelope states bit-equal."""
    sr, V, total = 48000.0, 70, 24 * 1024 + 77
    wl = synth.polyphony_bank(V, total, sr, seed=8, mean_on=0.1, mean_off=0.04)
    from zang_b200.events import BankTrigger
    off, spans = BankTrigger(wl["event_off"], wl["events"], total, 1024).spans(0, total)
    ev = wl["events"]
    pp = np.zeros(ev.size, dtype=L.param_dtypes[L.MOD_PULSE_OSC])
    pp["sample_rate"], pp["freq"], pp["color"] = sr, ev["freq"], 0.4
    dp = np.zeros(ev.size, dtype=L.param_dtypes[L.MOD_DECIMATOR])
    dp["sample_rate"], dp["fake_sample_rate"] = sr, 3000.0 + 500.0 * (np.arange(ev.size) % 7)
    fp = np.zeros(ev.size, dtype=L.param_dtypes[L.MOD_FILTER])
    fp["type"], fp["cutoff"], fp["res"] = L.FILTER_LOW_PASS, 0.15 + 0.05 * (np.arange(ev.size) % 5), 0.5
    ep = np.zeros(ev.size, dtype=L.param_dtypes[L.MOD_ENVELOPE])
    ep["sample_rate"], ep["attack"], ep["decay"], ep["release"] = sr, (L.CURVE_CUBED, 0.01), (L.CURVE_SQUARED, 0.08), (L.CURVE_LINEAR, 0.2)
    ep["sustain_volume"], ep["note_on"] = 0.6, ev["note_on"]
    results = []
    for mode, segments in ((L.PAINT_EXACT, 0), (L.PAINT_TOLERANCE, 6), (L.PAINT_TOLERANCE, 0)):
        b = E.GraphBuilder()
        b.zero(E.T(0))
        f_osc = b.call(L.MOD_PULSE_OSC, E.T(0))
        b.zero(E.T(1))
        f_dec = b.call(L.MOD_DECIMATOR, E.T(1), (E.T(0),))
        b.zero(E.T(2))
        f_flt = b.call(L.MOD_FILTER, E.T(2), (E.T(1),))
        b.zero(E.T(0))
        f_env = b.call(L.MOD_ENVELOPE, E.T(0))
        b.multiply(E.OUT(0), E.T(2), E.T(0))
        g = E.Graph(ctx, E.graph_config(V, b, 3), b)
        osc, dec, flt, env = E.PulseOsc(ctx, V), E.Decimator(ctx, V), E.Filter(ctx, V), E.Envelope(ctx, V)
        for f, m, rows in ((f_osc, osc, pp), (f_dec, dec, dp), (f_flt, flt, fp), (f_env, env, ep)):
            g.set_field(f, m.bank, rows)
        g.set_segments(segments)
        out = ctx.buf(total, V)
        ctx.zero(None, out)
        sp = ctx.spans(V, total, off, spans)
        ctx.set_paint_mode(mode)
        try:
            g.paint([sp], [out])
            info = g.segment_info()
            retries = g.warm_retries()
        finally:
            ctx.set_paint_mode(L.PAINT_EXACT)
        assert info["failed"] == 0 and retries == 0, (info, retries)
        assert (info["segments"] > 1) == (mode == L.PAINT_TOLERANCE), info
        if mode == L.PAINT_TOLERANCE:
            assert info["replay_frames"] >= int(np.ceil(sr / 3000.0)) + 2, info
        results.append((out.download(), osc.bank.get_state().tobytes() + dec.bank.get_state().tobytes() + env.bank.get_state().tobytes(),
                        flt.bank.get_state().view(np.float32).copy()))
        g.free()
    assert float(np.abs(results[0][0]).max()) > 0.1
    for r in results[1:]:
        assert_close(r[0], results[0][0], 1e-5, "Decimator + Filter graph in time segments vs one segment")
        assert r[1] == results[0][1], "carried PulseOsc / Decimator / Envelope state"
        assert np.abs(r[2] - results[0][2]).max() <= 1e-5, "carried Filter state"


@pytest.mark.parametrize("segments,warm", [(0, None), (5, None), (7, "48")], ids=["planned", "5seg", "7seg-short-window"])
def test_sfx_graph_filter_time_segments(ctx, segments, warm, monkeypatch):
    """A graph with a Filter in time segments (ZCU_PAINT_TOLERANCE): a segment skips the exact modules ahead to a window before
    its start (Noise by its GF(2) jump, Curve and Envelope in closed form), runs the whole graph silently over the window with
    the Filter starting from zero, and paints only if the window's homogeneous map has shrunk below 1e-8 on every voice of the
    warp; otherwise it lengthens the window and starts over (forced here with a 48-frame first guess).  Voices that paint
    nothing before the window keep their loaded state.  Two blocks, so the second starts from carried Filter states.
    Bar: 1e-5 of full scale against the per-module path in exact mode; Noise / Curve / Envelope states bit-equal."""
    if warm:
        monkeypatch.setenv("ZCU_GRAPH_WARM", warm)
    sr, jobs, job_frames, block = 48000.0, 70, 24576, 12288
    wl = synth.sfx_bank(jobs, job_frames, seed=5)
    ref = banks.SfxBank(ctx, wl, sr, job_frames, first_seed=9)
    unfused = _render(ctx, ref, job_frames, block, 3, fused=False)
    bank = banks.SfxBank(ctx, wl, sr, job_frames, first_seed=9)
    g = bank.build_graph()
    g.set_segments(segments)
    ctx.set_paint_mode(L.PAINT_TOLERANCE)
    try:
        got = _render(ctx, bank, job_frames, block, 3, fused=True)
        info = g.segment_info()
        retries = g.warm_retries()
    finally:
        ctx.set_paint_mode(L.PAINT_EXACT)
    assert info["failed"] == 0, info
    # the cutoff comes from a buffer: nothing is known about the decay, so segments only when asked for
    assert (info["segments"] > 1) == (segments != 0), info
    if warm is not None:
        assert retries > 0, "the 48-frame guess cannot have been enough"
    assert np.abs(unfused).max() > 0.05
    assert_close(got, unfused, 1e-5, "sfx graph in time segments vs per-module path")
    a, b = _states(bank), _states(ref)
    for k in a:
        if k == "filter":
            fa, fb = a[k].view(np.float32), b[k].view(np.float32)
            assert np.abs(fa - fb).max() <= 1e-5, "carried Filter state"
        else:
            assert a[k].tobytes() == b[k].tobytes(), f"carried state of {k}"
