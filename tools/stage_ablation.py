"""Profiling aid: time the fused render kernel with pipeline stages switched off (ZCU_RENDER_DEBUG_SKIP_MASK).
Results with a non-zero mask are numerically wrong on purpose; this only shows which stage sets the pace."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CODE = r'''
import sys, os
sys.path.insert(0, %r)
import numpy as np
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
print("mask", os.environ.get("ZCU_RENDER_DEBUG_SKIP_MASK", "0"), "render-kernel ms per 20 s job:", round(ms / 3, 2), "cycles/sample @1.965GHz:", round(ms / 3 * 1e-3 / total * 1.965e9, 1))
''' % ROOT
for mask in (0, 1, 2, 4, 8, 14, 13, 11, 7, 15):
    env = dict(os.environ, ZCU_RENDER_DEBUG_SKIP_MASK=str(mask))
    subprocess.run([sys.executable, "-c", CODE], env=env, check=True)
