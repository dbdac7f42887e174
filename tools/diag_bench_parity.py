"""Diagnostic: where does the SEGMENTS render of the bench job differ from the oracle inside the first segment?
Renders the first NV voices of the bench workload over the first segment in EXACT mode (64-voice sub-bank),
in SEGMENTS mode (whole job, voice window) and with the oracle, and prints the first differing samples."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle  # noqa: E402
import bench_parity  # noqa: E402
from zang_b200 import _lib as L  # noqa: E402
from zang_b200 import engine as E  # noqa: E402
from zang_b200 import synth  # noqa: E402

SR, BF, V, TOTAL = 48000.0, 1024, 4096, 2_880_000
NV = int(os.environ.get("NV", "64"))
N = int(os.environ.get("NFRAMES", str(105 * 1024)))

ctx = E.Context(0)
wl = synth.polyphony_bank(V, TOTAL, SR)
sub = bench_parity._sub(wl, 0, NV)
_, want, _ = oracle.render_nice(sub, N, SR, BF, want_mix=False, window=(0, N), nthreads=16)

# EXACT mode on the sub-bank
re_ = E.Render(ctx, NV, N, sub["color"], sub["gains"], channels=2, sample_rate=SR, buffer_frames=BF, mode=L.RENDER_EXACT)
ev_in = sub["events"][sub["events"]["frame"] < N] if False else None
# events beyond N frames are not allowed in a shorter job: rebuild offsets keeping only frames < N
off = sub["event_off"]
keep_off = [0]
keep = []
for v in range(NV):
    e = sub["events"][off[v]:off[v + 1]]
    e = e[e["frame"] < N]
    keep.append(e)
    keep_off.append(keep_off[-1] + e.size)
ev_cut = np.concatenate(keep)
off_cut = np.asarray(keep_off, dtype=np.uint32)
re_.set_events(off_cut, ev_cut)
oe = ctx.buf(N, NV)
re_.run_voices(0, oe)
got_exact = oe.download()

# SEGMENTS mode, whole job
rs = E.Render(ctx, V, TOTAL, wl["color"], wl["gains"], channels=2, sample_rate=SR, buffer_frames=BF, mode=L.RENDER_SEGMENTS)
rs.set_events(wl["event_off"], wl["events"])
mix = ctx.buf(TOTAL, 2)
rs.run(mix)
print("segment_info", rs.segment_info())
os_ = ctx.buf(N, NV)
rs.run_voice_window(0, 0, os_)
got_seg = os_.download()

# SEGMENTS mode on the cut sub-bank with one segment forced
r1 = E.Render(ctx, NV, N, sub["color"], sub["gains"], channels=2, sample_rate=SR, buffer_frames=BF, mode=L.RENDER_SEGMENTS)
r1.set_segments(1)
r1.set_events(off_cut, ev_cut)
o1 = ctx.buf(N, NV)
r1.run_voice_window(0, 0, o1)
got_seg1 = o1.download()


def report(name, got):
    d = got.view(np.uint32) != want.view(np.uint32)
    print(f"== {name}: {int(d.sum())} differing samples of {d.size}, max abs err {float(np.abs(got - want).max()):.3g}")
    if not d.any():
        return
    vs = np.nonzero(d.any(axis=0))[0]
    print("   voices with differences:", vs[:40], "count", vs.size)
    for v in vs[:4]:
        fr = np.nonzero(d[:, v])[0]
        f0 = int(fr[0])
        print(f"   voice {v}: {fr.size} frames differ, first {f0}, last {int(fr[-1])}; got {got[f0, v]!r} want {want[f0, v]!r} "
              f"(diff {got[f0, v] - want[f0, v]:.3g})")
        e = sub["events"][off[v]:off[v + 1]]
        near = e[(e["frame"] <= f0)][-3:]
        nxt = e[(e["frame"] > f0)][:2]
        print("     events before:", [tuple(x) for x in near])
        print("     events after :", [tuple(x) for x in nxt])
        # run structure of the differences
        brk = np.nonzero(np.diff(fr) > 1)[0]
        print(f"     difference runs: {brk.size + 1}; first run [{f0}, {int(fr[brk[0]]) if brk.size else int(fr[-1])}]")


report("EXACT (sub-bank)", got_exact)
report("SEGMENTS whole job, voice window", got_seg)
report("SEGMENTS sub-bank, 1 segment", got_seg1)
d2 = got_seg.view(np.uint32) != got_exact.view(np.uint32)
print("SEGMENTS vs EXACT differing samples:", int(d2.sum()))
