"""Profiling aid: per-stage busy fraction of the fused render kernel (ZCU_RENDER_DEBUG_CYCLES=1)."""
import os, sys
os.environ["ZCU_RENDER_DEBUG_CYCLES"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from zang_b200 import synth
from zang_b200.engine import Context, Render
V, secs = 4096, 20
total = secs * 48000
ctx = Context(0)
wl = synth.polyphony_bank(V, total)
r = Render(ctx, V, total, wl["color"], wl["gains"])
r.set_events(wl["event_off"], wl["events"])
mix = ctx.buf(total, 2)
r.run(mix); ctx.sync()
r.set_profiling(True)
for _ in range(3): r.run(mix)
ms, n = r.kernel_ms()
print("render-kernel ms per 20 s job:", round(ms / 3, 2), "cycles/sample @1.965GHz:", round(ms / 3 * 1e-3 / total * 1.965e9, 1))
r.free()
