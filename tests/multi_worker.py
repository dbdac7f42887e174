"""Worker of tests/test_gpu_multi.py: one process per GPU (spawned), no torch in the process — the host side a Zig caller
would write: zcu_ctx_create(rank) -> render the rank's voice shard -> zcu_nccl_init -> zcu_nccl_reduce_mix -> rank 0 reads
the summed mix back."""
import ctypes as C
import os
import sys

import numpy as np


def run(rank, world, nccl_id, voices, total, sr, mode, segments, out_queue):
    try:
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        for p in (root, os.path.join(root, "tests")):
            if p not in sys.path:
                sys.path.insert(0, p)
        from zang_b200 import _lib as L
        from zang_b200 import engine as E
        from zang_b200 import synth
        lib = L.lib
        ctx = E.Context(rank)
        wl_all = synth.polyphony_bank(voices, total, sr, seed=77, mean_on=0.05, mean_off=0.03)
        wl = synth.shard_voices(wl_all, rank, world)
        V = int(wl["color"].size)
        r = E.Render(ctx, V, total, wl["color"], wl["gains"], channels=2, sample_rate=sr, mode=mode)
        if mode == L.RENDER_SEGMENTS and segments:
            r.set_segments(segments)
        r.set_events(wl["event_off"], wl["events"])
        mix = ctx.buf(total, 2)
        ctx.zero(None, mix)
        r.run(mix)
        idb = (C.c_uint8 * 128).from_buffer_copy(nccl_id)
        rc = lib.zcu_nccl_init(ctx._h, world, rank, idb)
        assert rc == L.OK, (rc, lib.zcu_last_error(ctx._h))
        assert lib.zcu_nccl_init(ctx._h, world, rank, idb) == L.ERR_INVALID  # twice without finalize
        rc = lib.zcu_nccl_reduce_mix(ctx._h, mix._h, 0)
        assert rc == L.OK, (rc, lib.zcu_last_error(ctx._h))
        assert lib.zcu_nccl_reduce_mix(ctx._h, mix._h, world) == L.ERR_INVALID  # bad root
        got = mix.download() if rank == 0 else None
        assert lib.zcu_nccl_finalize(ctx._h) == L.OK
        assert lib.zcu_nccl_finalize(ctx._h) == L.OK  # idempotent
        r.free()
        ctx.close()
        out_queue.put((rank, "ok", got))
    except BaseException as e:  # noqa: BLE001  (report to the parent instead of dying silently)
        import traceback
        out_queue.put((rank, "error", traceback.format_exc() + repr(e)))
