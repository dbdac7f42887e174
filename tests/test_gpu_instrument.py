"""The polyphony-bank workload (BASELINE.json configs[1], scaled to what the oracle renders in seconds):
NiceInstrument (examples/modules.zig:189-248) driven by per-voice note streams through the Trigger.

Three implementations must agree:
  oracle     — per voice: ImpulseQueue -> Trigger -> NiceInstrument.paint, as example_polyphony.zig:61-92
  un-fused   — the drop-in module path: host Trigger -> span table -> zcu_zero / zcu_paint_pulse_osc / ...
  fused      — zcu_render_* (device-side Trigger replay, all module state in registers)
Per-voice output is BIT-EXACT in both device paths (no transcendental is evaluated per sample; the one
per-note cos/sqrt of Filter.cutoffFromFrequency is a host scalar in all three).  The stereo mix is held
to 1e-5 x sum|gain| because the reference's voice summation order is not reproduced.
"""
import numpy as np
import pytest

import oracle
from helpers import assert_bit_exact, assert_close
from zang_b200 import _lib as L
from zang_b200 import engine as E
from zang_b200 import synth

pytestmark = pytest.mark.gpu

SR = 48000.0


def small_workload(voices, total_frames, seed=123):
    # short notes so that attack/decay/sustain/release, retriggers and same-buffer events all occur
    wl = synth.polyphony_bank(voices, total_frames, SR, seed=seed, mean_on=0.05, mean_off=0.03)
    return wl


@pytest.mark.parametrize("voices,total", [(70, 8 * 1024), (33, 5 * 1024 + 300)])
def test_fused_voices_bit_exact(ctx, voices, total):
    wl = small_workload(voices, total)
    _, want, _ = oracle.render_nice(wl, total, SR, want_mix=False, window=(0, total))
    r = E.Render(ctx, voices, total, wl["color"], wl["gains"], channels=2, sample_rate=SR, block_buffers=3)
    r.set_events(wl["event_off"], wl["events"])
    vo = ctx.buf(total, voices)
    r.run_voices(0, vo)
    assert_bit_exact(vo.download(), want, "fused NiceInstrument voices")
    # a window that starts mid-render (state carried through blocks that are not returned)
    first, n = 2 * 1024 + 17, 1500
    _, want_w, _ = oracle.render_nice(wl, total, SR, want_mix=False, window=(first, n))
    vw = ctx.buf(n, voices)
    r.run_voices(first, vw)
    assert_bit_exact(vw.download(), want_w, "fused voices window")
    r.free()


@pytest.mark.parametrize("channels", [1, 2])
def test_fused_mix(ctx, channels):
    voices, total = 200, 6 * 1024 + 100
    wl = small_workload(voices, total, seed=7)
    gains = wl["gains"][:, :channels].copy()
    wl2 = dict(wl, gains=gains)
    want, _, _ = oracle.render_nice(wl2, total, SR, channels=channels)
    r = E.Render(ctx, voices, total, wl["color"], gains, channels=channels, sample_rate=SR, block_buffers=2)
    r.set_events(wl["event_off"], wl["events"])
    mix = ctx.buf(total, channels)
    r.run(mix)
    got = mix.download()
    assert np.abs(want).max() > 0.05
    assert_close(got, want, 1e-5 * float(np.abs(gains).sum(axis=0).max()), "fused mix")
    # accumulate semantics + determinism: a second run adds the identical render on top
    r.run(mix)
    assert_bit_exact(mix.download(), got + got, "second run must add the same bits")
    r.free()


SCAN_TOL = 1e-5  # BASELINE.json north_star: max abs error <= 1e-5 of full scale (a voice's full scale is 1.0)


@pytest.fixture(params=["scan", "segments"])
def tmode(request, monkeypatch):
    """The two time-parallel render modes.  SEGMENTS sizes its segments to the GPU, which on a small test job means one
    segment (= the exact render), so the tests force several segments per launch."""
    if request.param == "segments":
        monkeypatch.setenv("ZCU_RENDER_SEGMENTS_PER_LAUNCH", "5")
        return L.RENDER_SEGMENTS
    return L.RENDER_SCAN


@pytest.mark.parametrize("voices,total,bb", [(70, 8 * 1024, 3), (33, 5 * 1024 + 300, 0), (64, 40 * 1024, 16)])
def test_scan_mode_voices_within_tolerance(ctx, tmode, voices, total, bb):
    """ZCU_RENDER_SCAN (time-parallel chunks, zcu_render_scan.cuh): PulseOsc and Envelope start every chunk from an
    exact closed form, the Filter from a chained f64 2x2 scan, so per-voice output is held to the north-star
    tolerance instead of bit-exactness; where the filter has not yet run (before a voice's first note) it is exact."""
    wl = small_workload(voices, total, seed=31 + voices)
    _, want, _ = oracle.render_nice(wl, total, SR, want_mix=False, window=(0, total))
    r = E.Render(ctx, voices, total, wl["color"], wl["gains"], channels=2, sample_rate=SR, block_buffers=bb, mode=tmode)
    r.set_events(wl["event_off"], wl["events"])
    vo = ctx.buf(total, voices)
    r.run_voices(0, vo)
    got = vo.download()
    assert np.abs(want).max() > 0.1
    assert_close(got, want, SCAN_TOL, "scan-mode NiceInstrument voices")
    # the first mix buffer starts from the true (zero) filter state: bit-exact there
    assert_bit_exact(got[:1024], want[:1024], "scan mode, first chunk")
    first, n = 2 * 1024 + 17, 1500
    vw = ctx.buf(n, voices)
    r.run_voices(first, vw)
    assert_close(vw.download(), want[first:first + n], SCAN_TOL, "scan-mode voices window")
    r.free()


def test_segments_mode_with_one_segment_is_the_exact_render(ctx, monkeypatch):
    monkeypatch.setenv("ZCU_RENDER_SEGMENTS_PER_LAUNCH", "1")
    voices, total = 50, 7 * 1024 + 11
    wl = small_workload(voices, total, seed=5)
    _, want, _ = oracle.render_nice(wl, total, SR, want_mix=False, window=(0, total))
    r = E.Render(ctx, voices, total, wl["color"], wl["gains"], channels=2, sample_rate=SR, block_buffers=3, mode=L.RENDER_SEGMENTS)
    r.set_events(wl["event_off"], wl["events"])
    vo = ctx.buf(total, voices)
    r.run_voices(0, vo)
    assert_bit_exact(vo.download(), want, "segments mode, one segment per launch")
    r.free()


@pytest.mark.parametrize("nseg", ["1", "5"])
def test_segments_mode_with_contracted_filter(ctx, monkeypatch, nseg):
    """zcu_render_set_filter_contraction: the Filter recurrence with FMAs.  Per-voice output within the mode's tolerance (long
    sustained and slowly damped notes included: the rounding differences must not pile up), measurably not the exact
    bits any more, and the mix within its own tolerance.  Off again = the strict render.  Other modes refuse it."""
    monkeypatch.setenv("ZCU_RENDER_SEGMENTS_PER_LAUNCH", nseg)
    voices, total = 64, 40 * 1024
    wl = synth.polyphony_bank(voices, total, SR, seed=91, mean_on=0.3, mean_off=0.05)
    wl["events"]["freq"][::3] *= np.float32(0.125)  # every third event three octaves down: slow filters
    want_mix, want, _ = oracle.render_nice(wl, total, SR, want_mix=True, window=(0, total))
    r = E.Render(ctx, voices, total, wl["color"], wl["gains"], channels=2, sample_rate=SR, block_buffers=16, mode=L.RENDER_SEGMENTS)
    r.set_events(wl["event_off"], wl["events"])
    vo = ctx.buf(total, voices)
    r.run_voices(0, vo)
    strict = vo.download()
    r.set_filter_contraction(True)
    r.run_voices(0, vo)
    got = vo.download()
    assert np.abs(want).max() > 0.1
    assert_close(got, want, SCAN_TOL, "contracted Filter, voices")
    err = float(np.abs(got.astype(np.float64) - want).max())
    assert err < 2e-6, err  # a few ulp of the filter state, far inside the tolerance
    assert (got != strict).any()
    mix = ctx.buf(total, 2)
    r.run(mix)
    assert_close(mix.download(), want_mix, 1e-5 * float(np.abs(wl["gains"]).sum(axis=0).max()), "contracted Filter, mix")
    r.set_filter_contraction(False)
    r.run_voices(0, vo)
    assert_bit_exact(vo.download(), strict, "contraction off again")
    r.free()
    r2 = E.Render(ctx, voices, total, wl["color"], wl["gains"], channels=2, sample_rate=SR, block_buffers=16, mode=L.RENDER_EXACT)
    with pytest.raises(E.ZangCuError):
        r2.set_filter_contraction(True)
    r2.free()


def test_segments_mode_warm_up_is_adaptive(ctx, monkeypatch):
    """Low notes (cutoff = 8 x freq, resonance 0.7) damp slowly: the warm-up window must grow until its homogeneous
    map is <= 1e-8, and the result must still be within tolerance."""
    monkeypatch.setenv("ZCU_RENDER_SEGMENTS_PER_LAUNCH", "4")
    voices, total = 32, 48 * 1024
    wl = synth.polyphony_bank(voices, total, SR, seed=77, mean_on=0.3, mean_off=0.02)
    wl["events"]["freq"] *= np.float32(0.125)  # three octaves down: decay over several mix buffers
    _, want, _ = oracle.render_nice(wl, total, SR, want_mix=False, window=(0, total))
    r = E.Render(ctx, voices, total, wl["color"], wl["gains"], channels=2, sample_rate=SR, block_buffers=0, mode=L.RENDER_SEGMENTS)
    r.set_events(wl["event_off"], wl["events"])
    vo = ctx.buf(total, voices)
    r.run_voices(0, vo)
    assert_close(vo.download(), want, SCAN_TOL, "segments mode, slow filters")
    info = r.segment_info()
    assert info["segments"] == 4 and info["warm_max"] >= 4, info
    r.free()


def test_scan_mode_matches_exact_mode_mix(ctx, tmode):
    voices, total = 200, 6 * 1024 + 100
    wl = small_workload(voices, total, seed=7)
    want, _, _ = oracle.render_nice(wl, total, SR, channels=2)
    mixes = []
    for mode in (L.RENDER_EXACT, tmode):
        r = E.Render(ctx, voices, total, wl["color"], wl["gains"], channels=2, sample_rate=SR, block_buffers=2, mode=mode)
        r.set_events(wl["event_off"], wl["events"])
        mix = ctx.buf(total, 2)
        r.run(mix)
        mixes.append(mix.download())
        r.run(mix)
        assert_bit_exact(mix.download(), mixes[-1] + mixes[-1], "second run must add the same bits")
        r.free()
    tol = 1e-5 * float(np.abs(wl["gains"]).sum(axis=0).max())
    assert_close(mixes[1], want, tol, "scan-mode mix vs oracle")
    assert_close(mixes[1], mixes[0], tol, "scan-mode mix vs exact-mode mix")


@pytest.mark.parametrize("overwrite", [False, True])
def test_unfused_module_path_bit_exact(ctx, overwrite):
    """The drop-in path: every module called through its own C-ABI entry point on device buffers.  overwrite: the body's
    `zero; paint` pairs as single calls in ZCU_OUT_OVERWRITE mode (5 calls instead of 8), temps pre-filled with garbage."""
    voices, nbuf = 40, 6
    total = nbuf * 1024
    wl = small_workload(voices, total, seed=99)
    _, want, _ = oracle.render_nice(wl, total, SR, want_mix=False, window=(0, total))
    instr_params = np.zeros(wl["events"].size, dtype=E.NiceInstrumentBank.Params)
    instr_params["sample_rate"] = SR
    instr_params["freq"] = wl["events"]["freq"]
    instr_params["note_on"] = wl["events"]["note_on"]
    # NiceInstrument.color is per instance; the un-fused bank takes one colour, so paint voice groups by colour
    color = np.float32(0.3)
    wl_c = dict(wl, color=np.full(voices, color, dtype=np.float32))
    _, want, _ = oracle.render_nice(wl_c, total, SR, want_mix=False, window=(0, total))
    bank = E.NiceInstrumentBank(ctx, voices, color)
    out, t0, t1 = ctx.buf(total, voices), ctx.buf(total, voices), ctx.buf(total, voices)
    # two blocks of 3 buffers: module state and Trigger state carried across the block boundary
    half = 3 * 1024
    for blk in range(2):
        off, spans = synth.spans_from_events(wl["event_off"], wl["events"], total, 1024, first_frame=blk * half, num_frames=half)
        sp = ctx.spans(voices, half, off, spans)
        junk = np.random.default_rng(blk).standard_normal((half, voices)).astype(np.float32)
        o, a, b = ctx.buf(half, voices), ctx.buf(half, voices, junk), ctx.buf(half, voices, junk * 3)
        bank.paint(sp, [o], [a, b], instr_params, overwrite=overwrite)
        assert ctx.paint_output() == L.OUT_ACCUMULATE
        got = o.download()
        assert_bit_exact(got, want[blk * half:(blk + 1) * half], f"un-fused NiceInstrument block {blk}")
        for x in (sp, o, a, b):
            x.free()


def test_render_config_errors(ctx):
    import ctypes as C
    from zang_b200._lib import lib, ptr
    cfg = L.RenderConfig(L.INSTR_NICE, 8, 1024, 4096, 0, 3, SR, 0, L.RENDER_EXACT)
    col, g = np.zeros(8, np.float32), np.zeros((8, 3), np.float32)
    h = C.c_void_p()
    assert lib.zcu_render_create(ctx._h, C.byref(cfg), ptr(col), ptr(g), C.byref(h)) == L.ERR_INVALID
    cfg.channels = 2
    assert lib.zcu_render_create(ctx._h, C.byref(cfg), ptr(col), ptr(g), C.byref(h)) == L.OK
    mix = ctx.buf(4096, 2)
    assert lib.zcu_render_run(ctx._h, h, mix._h) == L.ERR_INVALID  # no events yet
    lib.zcu_render_free(h)


def test_set_events_rejects_malformed_tables(ctx):
    """zcu_render_set_events validates what the device kernels will index with: offsets monotone from 0,
    a voice's frames chronological and inside the render.  A bad table is ZCU_ERR_INVALID, not a GPU fault."""
    voices, total = 8, 4 * 1024
    wl = small_workload(voices, total, seed=3)
    r = E.Render(ctx, voices, total, wl["color"], wl["gains"], channels=2, sample_rate=SR, mode=L.RENDER_SEGMENTS)
    off, ev = wl["event_off"].copy(), wl["events"].copy()
    bad_off = off.copy()
    bad_off[3], bad_off[4] = off[4], off[3]  # not monotone
    with pytest.raises(E.ZangCuError):
        r.set_events(bad_off, ev)
    bad_ev = ev.copy()
    bad_ev["frame"][off[2]] = total  # outside the render
    with pytest.raises(E.ZangCuError):
        r.set_events(off, bad_ev)
    bad_ev = ev.copy()
    k = int(off[5])
    assert off[6] - off[5] >= 2
    bad_ev["frame"][k], bad_ev["frame"][k + 1] = ev["frame"][k + 1] + 1, ev["frame"][k]  # out of order inside a voice
    with pytest.raises(E.ZangCuError):
        r.set_events(off, bad_ev)
    shifted = off.copy()
    shifted[0] = 1
    with pytest.raises(E.ZangCuError):
        r.set_events(shifted, ev)
    # and the object still works after the rejections
    r.set_events(off, ev)
    _, want, _ = oracle.render_nice(wl, total, SR, want_mix=False, window=(0, total))
    vo = ctx.buf(total, voices)
    r.run_voices(0, vo)
    assert_close(vo.download(), want, SCAN_TOL, "render after rejected tables")
    r.free()
