"""The benchmark job itself against the oracle: BASELINE.json configs[1] exactly as bench.py builds it
(`synth.polyphony_bank(4096, 2_880_000, 48000.0)`, ZCU_RENDER_SEGMENTS, default segment sizing), checked per
SURVEY.md §8(d) by tests/bench_parity.py.  Reference graph: examples/example_polyphony.zig:61-92 x
examples/modules.zig:189-248."""
import os

import numpy as np
import pytest

import bench_parity
from zang_b200 import _lib as L
from zang_b200 import engine as E
from zang_b200 import synth

pytestmark = pytest.mark.gpu

SR, BF = 48000.0, 1024
V, TOTAL = 4096, 2_880_000


@pytest.fixture(scope="module")
def bench_job(ctx):
    assert "ZCU_RENDER_SEGMENTS_PER_LAUNCH" not in os.environ, "the bench configuration uses the default segment sizing"
    wl = synth.polyphony_bank(V, TOTAL, SR)
    r = E.Render(ctx, V, TOTAL, wl["color"], wl["gains"], channels=2, sample_rate=SR, buffer_frames=BF, mode=L.RENDER_SEGMENTS)
    r.set_events(wl["event_off"], wl["events"])
    mix = ctx.buf(TOTAL, 2)
    r.run(mix)
    got = mix.download()
    yield ctx, r, wl, got, mix
    mix.free()
    r.free()


def test_bench_segmentation_is_the_one_bench_prints(bench_job):
    ctx, r, wl, got, mix = bench_job
    info = r.segment_info()
    # 148 SMs x 6 resident CTAs x 4 warps / 128 voice groups = 27 segments of ceil(2813 / 27) = 105 mix buffers
    assert info["segments"] * info["buffers_per_segment"] >= -(-TOTAL // BF)
    assert info["segments"] >= 8, info          # the job really is cut in time (one segment would be the exact render)
    assert 1 <= info["warm_max"] <= 8, info     # cutoff = 8 x freq, res 0.7: damped within a few buffers
    assert r.launches_per_run == 2              # one k_nice_seg_render + one k_partial_reduce for the whole job


def test_bench_job_voices_and_mix_against_the_oracle(bench_job):
    ctx, r, wl, got, mix = bench_job
    nthreads = max(1, min(32, len(os.sched_getaffinity(0))))
    res = bench_parity.check_bench_job(ctx, r, wl, TOTAL, SR, BF, mix_got=got, nthreads=nthreads)
    print("bench-config parity:", res)
    assert res["ok"] and res["windows"] >= 64 + res["segment_starts_probed"]
    assert res["segment_starts_probed"] == res["segments"] - 1
    assert res["bit_exact_voice_samples"] > 0


def test_bench_job_is_deterministic(bench_job):
    ctx, r, wl, got, mix = bench_job
    r.run(mix)  # accumulates: the second run must add the identical bits
    twice = mix.download()
    assert np.array_equal(twice.view(np.uint32), (got + got).view(np.uint32))
