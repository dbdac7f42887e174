"""Per-call device time of the un-fused NiceInstrument chain (examples/modules.zig:220-246) on the configs[1] bank:
which of the nine C-ABI calls of one block the time goes to.   python tools/nice_calls.py [--exact]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from zang_b200 import _lib as L  # noqa: E402
from zang_b200 import engine as E, synth  # noqa: E402
from zang_b200.events import BankTrigger  # noqa: E402

ctx = E.Context(0)
ctx.set_paint_mode(L.PAINT_EXACT if "--exact" in sys.argv else L.PAINT_TOLERANCE)
V, frames, sr = 4096, 64 * 1024, 48000.0
if "--first-block" in sys.argv:  # the job's first 64 buffers (voices start one by one: round 1's measurement)
    wl = synth.polyphony_bank(V, frames, sr)
    trig = BankTrigger(wl["event_off"], wl["events"], frames, 1024)
    prime = None
else:  # a steady-state block: buffers [1024, 1088) of the 60 s job, the preceding 64 buffers painted first
    wl = synth.window_bank(synth.polyphony_bank(V, 2_880_000, sr), 960 * 1024, 1088 * 1024)
    trig = BankTrigger(wl["event_off"], wl["events"], 2 * frames, 1024)
    prime = trig.spans(0, frames)
bank = E.NiceInstrumentBank(ctx, V, 0.3)
off, spans = trig.spans(frames if prime is not None else 0, frames)
sp = ctx.spans(V, frames, off, spans)
ev = wl["events"]
rows = np.zeros(ev.size, dtype=E.NiceInstrumentBank.Params)
rows["sample_rate"], rows["freq"], rows["note_on"] = sr, ev["freq"], ev["note_on"]
pp, fp, ep = bank.derive(rows)
out, t0, t1 = ctx.buf(frames, V), ctx.buf(frames, V), ctx.buf(frames, V)
if prime is not None:
    sp0 = ctx.spans(V, frames, *prime)
    bank.paint(sp0, [out], [t0, t1], rows)
    sp0.free()
print(f"spans: {spans.size} ({spans.size / V:.1f} per voice), covered fraction {float((spans['end'] - spans['start']).sum()) / (V * frames):.3f}")
calls = [("zero t0", 4, lambda: ctx.zero(sp, t0)), ("PulseOsc", 8, lambda: bank.osc.paint(sp, [t0], [], pp)),
         ("multiplyWithScalar", 8, lambda: ctx.multiplyWithScalar(sp, t0, 0.5)), ("zero t1", 4, lambda: ctx.zero(sp, t1)),
         ("Filter", 12, lambda: bank.flt.paint(sp, [t1], [], fp, input=t0)), ("zero t0", 4, lambda: ctx.zero(sp, t0)),
         ("Envelope", 8, lambda: bank.env.paint(sp, [t0], [], ep)), ("multiply", 16, lambda: ctx.multiply(sp, out, t0, t1))]
tot = 0.0
for name, bps, fn in calls:
    best = 1e9
    for rep in range(3):
        ctx.timer_start()
        fn()
        best = min(best, ctx.timer_stop_ms())
    tot += best
    import ctypes as C
    st = (C.c_uint32 * 4)()
    L.lib.zcu_ctx_paint_warm_stats(ctx._h, st)
    extra = f"   warm-up: {st[0] // 3} CTAs/launch, mean window {st[1] / max(st[0], 1):.0f} frames, max {st[2]}, chained {st[3] // 3}/launch" if st[0] else ""
    print(f"{name:20s} {best:8.3f} ms  {V * frames * bps / best / 1e6:8.1f} GB/s (full-coverage bytes){extra}")
print(f"sum {tot:.3f} ms = {V * frames * 64 / tot / 1e6:.0f} GB/s of the reference chain's 64 B per voice-sample")
# the same body with every `zero; paint` pair as ONE call (ZCU_OUT_OVERWRITE): 5 calls
ctx.set_paint_output(L.OUT_OVERWRITE)
calls2 = [("PulseOsc (overwrite)", lambda: bank.osc.paint(sp, [t0], [], pp)), ("multiplyWithScalar", lambda: ctx.multiplyWithScalar(sp, t0, 0.5)),
          ("Filter (overwrite)", lambda: bank.flt.paint(sp, [t1], [], fp, input=t0)), ("Envelope (overwrite)", lambda: bank.env.paint(sp, [t0], [], ep)),
          ("multiply", lambda: ctx.multiply(sp, out, t0, t1))]
tot2 = 0.0
for name, fn in calls2:
    best = 1e9
    for rep in range(3):
        ctx.timer_start()
        fn()
        best = min(best, ctx.timer_stop_ms())
    tot2 += best
    print(f"{name:22s} {best:8.3f} ms")
ctx.set_paint_output(L.OUT_ACCUMULATE)
print(f"sum {tot2:.3f} ms = {V * frames * 64 / tot2 / 1e6:.0f} GB/s of the reference chain's 64 B per voice-sample")
