"""Multi-GPU correctness on real GPUs: voices sharded over the GPUs of one box, one process per GPU, the
partial stereo mixes summed by `zcu_nccl_unique_id / zcu_nccl_init / zcu_nccl_reduce_mix / zcu_nccl_finalize` — the C-ABI
collective INTEGRATION.md §4 gives a non-Python host.  Rank 0's reduced mix is compared with the un-sharded render on one
GPU and with the oracle.  Skipped on a box with fewer than 2 GPUs (run with `gpurun --gpus 2`)."""
import ctypes as C
import multiprocessing as mp

import numpy as np
import pytest

import oracle
from helpers import assert_close
from zang_b200 import _lib as L
from zang_b200 import engine as E
from zang_b200 import synth

pytestmark = pytest.mark.gpu

SR = 48000.0


def _spawn(world, voices, total, mode, segments):
    import multi_worker
    idb = (C.c_uint8 * 128)()
    assert L.lib.zcu_nccl_unique_id(idb) == L.OK
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    procs = [mpc.Process(target=multi_worker.run, args=(r, world, bytes(idb), voices, total, SR, mode, segments, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = {}
    for _ in range(world):
        rank, status, payload = q.get(timeout=300)
        assert status == "ok", f"rank {rank}: {payload}"
        results[rank] = payload
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return results[0]


@pytest.mark.parametrize("mode,segments", [(L.RENDER_EXACT, 0), (L.RENDER_SEGMENTS, 3)])
def test_sharded_render_reduced_over_nccl_matches_one_gpu_and_the_oracle(ctx, mode, segments):
    world = min(4, L.lib.zcu_device_count())
    if world < 2:
        pytest.skip("needs at least 2 GPUs")
    voices, total = 203, 12 * 1024 + 100  # an uneven shard split, a partial last buffer
    got = _spawn(world, voices, total, mode, segments)
    wl = synth.polyphony_bank(voices, total, SR, seed=77, mean_on=0.05, mean_off=0.03)
    tol = 1e-5 * float(np.abs(wl["gains"]).sum(axis=0).max())
    want, _, _ = oracle.render_nice(wl, total, SR, channels=2)
    assert np.abs(want).max() > 0.05
    assert_close(got, want, tol, f"{world}-GPU reduced mix vs oracle")
    r = E.Render(ctx, voices, total, wl["color"], wl["gains"], channels=2, sample_rate=SR, mode=mode)
    if segments:
        r.set_segments(segments)
    r.set_events(wl["event_off"], wl["events"])
    mix = ctx.buf(total, 2)
    r.run(mix)
    assert_close(got, mix.download(), tol, f"{world}-GPU reduced mix vs the un-sharded render")
    mix.free()
    r.free()
