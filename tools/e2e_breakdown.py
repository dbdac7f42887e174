"""Where the end-to-end leg of bench.py spends its time beyond the render (configs[1], 4096 voices x 60 s).

Wall-clock of each piece with a stream synchronize on both sides: zcu_render_set_events (host expansion of the note events +
H2D + the per-voice prepass kernel) for several host thread counts, the bare H2D of the same bytes, the render, the D2H of the mix.
  python tools/e2e_breakdown.py [--out gpurun_out/e2e_breakdown.json]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from zang_b200 import _lib as L  # noqa: E402
from zang_b200 import synth  # noqa: E402
from zang_b200.engine import Context, Render  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    SR, V, frames = 48000.0, 4096, 60 * 48000
    stream = torch.cuda.current_stream()
    ctx = Context(0, stream=stream.cuda_stream)
    wl = synth.polyphony_bank(V, frames, SR)
    n = int(wl["event_off"][-1])
    ev_pinned = torch.empty(n * L.note_event_dtype.itemsize, dtype=torch.uint8).pin_memory()
    ev = ev_pinned.numpy().view(L.note_event_dtype)
    ev[:] = wl["events"]
    off = np.ascontiguousarray(wl["event_off"], dtype=np.uint32)
    mix_t = torch.zeros((frames, 2), dtype=torch.float32, device="cuda")
    mix_host = torch.empty((frames, 2), dtype=torch.float32).pin_memory()
    mix = ctx.wrap(mix_t.data_ptr(), frames, 2)
    r = Render(ctx, V, frames, wl["color"], wl["gains"], channels=2, sample_rate=SR, buffer_frames=1024, mode=L.RENDER_SEGMENTS)
    r.set_events(off, ev)
    staged = torch.empty(n * 32, dtype=torch.uint8).pin_memory()
    dev = torch.empty(n * 32, dtype=torch.uint8, device="cuda")

    def wall(fn):
        best = 1e30
        for _ in range(args.reps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            best = min(best, (time.perf_counter() - t0) * 1e3)
        return best

    res = {"events": n, "voices": V, "frames": frames}
    for th in (1, 2, 4, 8):
        os.environ["ZCU_HOST_THREADS"] = str(th)
        res[f"set_events_ms_threads{th}"] = wall(lambda: r.set_events(off, ev, sync=False))
    os.environ.pop("ZCU_HOST_THREADS")
    res["set_events_ms_default"] = wall(lambda: r.set_events(off, ev, sync=False))
    res["h2d_32B_per_event_ms"] = wall(lambda: dev.copy_(staged, non_blocking=True))
    res["zero_mix_ms"] = wall(lambda: mix_t.zero_())
    res["render_ms"] = wall(lambda: r.run(mix))
    res["d2h_mix_ms"] = wall(lambda: mix_host.copy_(mix_t, non_blocking=True))

    def whole():
        r.set_events(off, ev, sync=False)
        mix_t.zero_()
        r.run(mix)
        mix_host.copy_(mix_t, non_blocking=True)
    res["whole_ms"] = wall(whole)
    for k, v in res.items():
        print(f"{k:32s} {v:10.3f}" if isinstance(v, float) else f"{k:32s} {v}")
    if args.out:
        json.dump(res, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
