"""Parity gate of the BENCHED kernel at the BENCHED configuration (VERDICT r01 "Next round" #1).

`bench.py` times `zcu_render_run` in ZCU_RENDER_SEGMENTS mode on 4096 voices x 2 880 000 frames.  This module compares
that very job with the CPU oracle (checker only):

  (a) the first 64 voices over the first 8 s, per voice, un-mixed;
  (b) one two-buffer window across EVERY time-segment start of the launch (the Filter state there comes from a warm-up
      window, everything else from closed forms), on a random 32-voice group each;
  (c) 64 random (32-voice group, mix buffer) windows anywhere in the job;
  (d) the whole stereo mix.

Per-voice bar: 1e-5 of full scale (a voice's full scale is 1.0); frames of the launch's first segment start from the
true state and must be bit-exact.  Mix bar: 1e-5 x sum|gain|.  Voice windows go through
`zcu_render_run_voice_window`, which renders with the launches and segments of the whole job.

Used by tests/test_gpu_bench_config.py and, untimed, by bench.py (the `parity` object of its JSON line).
"""
import time

import numpy as np

import oracle
from zang_b200 import synth

VOICE_TOL = 1e-5


def _sub(wl, a, b):
    off = wl["event_off"]
    return {"event_off": (off[a:b + 1] - off[a]).astype(np.uint32), "events": wl["events"][off[a]:off[b]], "color": wl["color"][a:b],
            "gains": wl["gains"][a:b]}


def _oracle_window(wl, a, b, first, n, total_frames, sr, bf, nthreads):
    """Oracle output of voices [a, b) over frames [first, first + n): the voices are rendered from frame 0."""
    end = min(total_frames, -(-(first + n) // bf) * bf)
    _, vo, _ = oracle.render_nice(_sub(wl, a, b), end, sr, bf, want_mix=False, window=(first, n), nthreads=max(1, min(nthreads, b - a)))
    return vo


def check_bench_job(ctx, r, wl, total_frames, sr, bf, mix_got=None, mix_wl=None, mix_frames=None, nthreads=8, first_voices=64,
                    first_seconds=8.0, n_random=64, seed=20260711):
    """r: a Render in SEGMENTS mode with wl's events set, already run once (segment_info describes the benched launch).
    mix_got: [frames][C] host array of the job's mix (optional); mix_wl: the workload whose voices are in that mix
    (default wl); mix_frames: how many leading frames of the mix to check (default all).
    Returns a dict for the bench line; raises AssertionError on a parity failure."""
    t_begin = time.time()
    V = int(wl["color"].size)
    info = r.segment_info()
    nseg, bps = int(info["segments"]), int(info["buffers_per_segment"])
    seg_frames = bps * bf
    rng = np.random.default_rng(seed)
    ngroups = (V + 31) // 32
    res = {"segments": nseg, "buffers_per_segment": bps, "warm_max": int(info["warm_max"]), "voice_tol": VOICE_TOL}
    max_err, nwin, exact_frames, checked, silent = 0.0, 0, 0, 0, 0

    def probe(a, b, first, n):
        nonlocal max_err, nwin, exact_frames, checked, silent
        want = _oracle_window(wl, a, b, first, n, total_frames, sr, bf, nthreads)
        out = ctx.buf(n, b - a)
        r.run_voice_window(a, first, out)
        got = out.download()
        out.free()
        err = float(np.abs(got - want).max()) if got.size else 0.0
        assert np.isfinite(got).all(), f"non-finite output in window voices [{a},{b}) frames [{first},{first + n})"
        assert err <= VOICE_TOL, f"voices [{a},{b}) frames [{first},{first + n}): max abs err {err:.3g} > {VOICE_TOL}"
        # the launch's first segment starts from the true state: bit-exact there
        ex = max(0, min(first + n, seg_frames) - first)
        if ex:
            assert np.array_equal(got[:ex].view(np.uint32), want[:ex].view(np.uint32)), \
                f"first segment not bit-exact: voices [{a},{b}) frames [{first},{first + ex})"
            exact_frames += ex * (b - a)
        silent += int(float(np.abs(want).max()) == 0.0)
        max_err = max(max_err, err)
        nwin += 1
        checked += got.size
        return err

    # (a) first voices x first seconds
    fv = min(first_voices, V)
    fn = min(total_frames, int(first_seconds * sr) // bf * bf)
    if fv and fn:
        probe(0, fv, 0, fn)
    # (b) across every segment start
    straddle_err = 0.0
    for s in range(1, nseg):
        Fs = s * seg_frames
        if Fs >= total_frames:
            break
        g = int(rng.integers(0, ngroups))
        a, b = g * 32, min(V, g * 32 + 32)
        first = Fs - bf
        n = min(2 * bf, total_frames - first)
        straddle_err = max(straddle_err, probe(a, b, first, n))
    # (c) random (voice group, buffer) windows
    nbuf = (total_frames + bf - 1) // bf
    for _ in range(n_random):
        g = int(rng.integers(0, ngroups))
        a, b = g * 32, min(V, g * 32 + 32)
        first = int(rng.integers(0, nbuf)) * bf
        probe(a, b, first, min(bf, total_frames - first))
    res.update({"windows": nwin, "voice_samples_checked": int(checked), "max_abs_err_voices": max_err,
                "max_abs_err_at_segment_starts": straddle_err, "segment_starts_probed": max(0, min(nseg, -(-total_frames // seg_frames)) - 1),
                "bit_exact_voice_samples": int(exact_frames)})
    assert silent <= max(2, nwin // 16), f"{silent} of {nwin} oracle windows are silent: the probe is not looking at sound"
    # (d) the mix
    if mix_got is not None:
        mwl = wl if mix_wl is None else mix_wl
        mf = total_frames if mix_frames is None else min(total_frames, -(-int(mix_frames) // bf) * bf)
        want, _, secs = oracle.render_nice(mwl, mf, sr, bf, channels=mix_got.shape[1], nthreads=nthreads)
        tol = VOICE_TOL * float(np.abs(mwl["gains"]).sum(axis=0).max())
        merr = float(np.abs(mix_got[:mf] - want).max())
        assert float(np.abs(want).max()) > 0.05, "oracle mix is silent"
        assert merr <= tol, f"mix max abs err {merr:.3g} > {tol:.3g}"
        res.update({"mix_max_abs_err": merr, "mix_tol": tol, "mix_frames_checked": int(mf), "mix_voices": int(mwl["color"].size),
                    "oracle_mix_seconds": secs})
    res["ok"] = True
    res["seconds"] = time.time() - t_begin
    return res
