"""The N > 1 path on CPU (gloo, world_size 2): voices shard over ranks with no data-path exchange, the
per-rank partial mixes are summed with one reduce to rank 0.  The per-rank render is done by
the oracle here (no GPU in this container); what is under test is the sharding of the workload, the reduce
plumbing bench.py uses, and that shard-then-sum equals the un-sharded render within the mix tolerance."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, voices, total, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    import oracle
    from zang_b200 import synth
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    wl_all = synth.polyphony_bank(voices, total, 48000.0, seed=5, mean_on=0.05, mean_off=0.03)
    wl = synth.shard_voices(wl_all, rank, world)
    a, b = wl["voice_range"]
    assert wl["event_off"][0] == 0 and wl["event_off"][-1] == wl["events"].size and wl["color"].size == b - a
    mix, _, _ = oracle.render_nice(wl, total)
    t = torch.from_numpy(mix)
    dist.reduce(t, dst=0)  # the one collective of the path
    if rank == 0:
        q.put(t.numpy().copy())
    dist.barrier()
    dist.destroy_process_group()


def test_shard_and_reduce_equals_unsharded():
    import torch.multiprocessing as mp
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle
    from zang_b200 import synth
    voices, total, world = 37, 4 * 1024 + 100, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, voices, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    wl_all = synth.polyphony_bank(voices, total, 48000.0, seed=5, mean_on=0.05, mean_off=0.03)
    want, _, _ = oracle.render_nice(wl_all, total)
    tol = 1e-5 * float(np.abs(wl_all["gains"]).sum(axis=0).max())
    assert np.abs(want).max() > 0.01
    assert float(np.abs(got - want).max()) <= tol


def test_shards_partition_the_bank():
    from zang_b200 import synth
    wl = synth.polyphony_bank(50, 20000, 48000.0, seed=9)
    for world in (1, 2, 4, 8):
        seen_events, seen_voices = 0, 0
        for r in range(world):
            s = synth.shard_voices(wl, r, world)
            a, b = s["voice_range"]
            seen_voices += b - a
            seen_events += s["events"].size
            assert np.array_equal(s["events"], wl["events"][wl["event_off"][a]:wl["event_off"][b]])
            assert np.array_equal(s["gains"], wl["gains"][a:b])
        assert seen_voices == 50 and seen_events == wl["events"].size


def test_workload_generator_invariants():
    from zang_b200 import synth
    wl = synth.polyphony_bank(64, 5 * 48000, 48000.0)
    ev, off = wl["events"], wl["event_off"]
    for v in range(64):
        e = ev[off[v]:off[v + 1]]
        assert np.all(np.diff(e["frame"].astype(np.int64)) >= 0)          # chronological (ImpulseQueue.push drops otherwise)
        assert np.all(e["note_on"][0::2] == 1) and np.all(e["note_on"][1::2] == 0)  # on/off alternate
        assert np.array_equal(e["note_id"][0::2][:e["note_id"][1::2].size], e["note_id"][1::2])  # off carries its note's id
        assert np.all(np.diff(e["note_id"][0::2].astype(np.int64)) > 0)   # fresh id per note-on (Envelope.zig:45)
        per_buffer = np.bincount(e["frame"] // 1024)
        assert per_buffer.max() <= 32                                      # ImpulseQueue capacity (notes.zig:73-74)
    assert wl["color"].min() >= 0.1 and wl["color"].max() <= 0.9
