"""Throughput of the instrument banks of BASELINE.json configs[1..4] through the DROP-IN module path (every line of the
reference-style paint body = one C-ABI call over all voices), at the configs' voice counts, one block of mix buffers.

Reported per config: device time of the module chain for the block (CUDA events on the context stream, span tables already
uploaded), voice-samples/s, and the algorithmic HBM GB/s (SURVEY.md §8d bytes per voice-sample of the un-fused chain);
plus the host time to build and upload the block's span tables (Trigger replay in libzangcu).

  python tools/bank_bench.py [--out gpurun_out/bank_bench.json]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from zang_b200 import _lib as L  # noqa: E402
from zang_b200 import banks, synth  # noqa: E402
from zang_b200 import engine as E  # noqa: E402


def run(ctx, name, bank, frames, ntemps, bytes_per_vs, peak, ins=(), reps=3):
    V = bank.voices
    out = ctx.buf(frames, V)
    temps = [ctx.buf(frames, V) for _ in range(ntemps)]
    t_host = time.perf_counter()
    prepared = bank.prepare_block(0, frames)
    ctx.sync()
    t_host = time.perf_counter() - t_host
    best = 1e30
    launches0 = ctx.launch_count
    for rep in range(reps + 1):
        ctx.zero(None, out)
        ctx.timer_start()
        bank.paint_block(0, frames, out, temps, *ins, prepared=prepared)
        ms = ctx.timer_stop_ms()
        if rep > 0:
            best = min(best, ms)
    launches = (ctx.launch_count - launches0) // (reps + 1)
    vs = V * frames / (best * 1e-3)
    gbs = vs * bytes_per_vs / 1e9
    res = dict(config=name, voices=V, frames=frames, ms=best, voice_samples_per_s=vs, algorithmic_bytes_per_voice_sample=bytes_per_vs,
               gbs=gbs, frac_of_peak=gbs / peak, host_span_tables_ms=t_host * 1e3, launches_per_block=launches)
    print(f"{name:34s} V={V:6d} frames={frames:6d} {best:9.3f} ms  {vs:10.3e} vs/s  {gbs:8.1f} GB/s  {gbs / peak:5.3f} of peak   "
          f"(host span tables {t_host * 1e3:7.1f} ms, {launches} launches)", flush=True)
    for b in [out] + temps + list(prepared):
        b.free()
    return res


def run_fused(ctx, name, bank, frames, bytes_per_vs, peak, ins=(), reps=3, gains=None, wl=None):
    """The same bank body as ONE fused graph kernel (zcu_graph_*): per-voice output buffer (BUFFERS) and stereo mix (MIX).
    With wl: the block's span tables come from the device-side Trigger replay (zcu_event_bank_*); device_span_tables_ms is
    its device time, host_span_call_ms the host time of the call that enqueues it."""
    V = bank.voices
    res = {}
    if wl is not None:
        bank.use_device_spans(wl)
        bank.prepare_block(0, frames)  # warm-up (allocations, static tables)
        bank.dev_trigger.reset()
        ctx.sync()
        t0 = time.perf_counter()
        ctx.timer_start()
        prepared = bank.prepare_block(0, frames)
        t_call = time.perf_counter() - t0
        dev_ms = ctx.timer_stop_ms()
        res["device_span_tables_ms"] = dev_ms
        res["host_span_call_ms"] = t_call * 1e3
        print(f"{name:34s} device span tables: {dev_ms:7.3f} ms on the GPU, {t_call * 1e3:7.3f} ms of host time to enqueue", flush=True)
    else:
        prepared = bank.prepare_block(0, frames)
    for mode, mname in ((L.GRAPH_OUT_BUFFERS, "fused_buffers"), (L.GRAPH_OUT_MIX, "fused_mix")):
        if mode == L.GRAPH_OUT_MIX and gains is None:
            continue
        t0 = time.perf_counter()
        bank.build_graph(output_mode=mode, channels=2, gains=gains if mode == L.GRAPH_OUT_MIX else None)
        t_build = time.perf_counter() - t0
        out = ctx.buf(frames, V if mode == L.GRAPH_OUT_BUFFERS else 2)
        best = 1e30
        for rep in range(reps + 1):
            ctx.zero(None, out)
            ctx.timer_start()
            bank.paint_block_fused(0, frames, out, *ins, prepared=prepared)
            ms = ctx.timer_stop_ms()
            if rep > 0:
                best = min(best, ms)
        vs = V * frames / (best * 1e-3)
        res[mname] = dict(ms=best, voice_samples_per_s=vs, graph_build_s=t_build, equivalent_algorithmic_gbs=vs * bytes_per_vs / 1e9,
                          equivalent_frac_of_peak=vs * bytes_per_vs / 1e9 / peak)
        print(f"{name:34s} {mname:14s} {best:9.3f} ms  {vs:10.3e} vs/s  (graph build {t_build:5.2f} s; un-fused-equivalent {vs * bytes_per_vs / 1e9:8.1f} GB/s)", flush=True)
        if mode == L.GRAPH_OUT_MIX and wl is not None:
            # end to end for one block, wall clock: device-side Trigger replay -> fused kernel -> partial reduce -> D2H of the mix
            e2e = 1e30
            for rep in range(3):
                bank.dev_trigger.reset()
                ctx.zero(None, out)
                ctx.sync()
                t0 = time.perf_counter()
                tabs = bank.prepare_block(0, frames)
                bank.paint_block_fused(0, frames, out, *ins, prepared=tabs)
                host_mix = out.download()
                e2e = min(e2e, time.perf_counter() - t0)
                for t in tabs:
                    t.free()
            res[mname]["e2e_block_ms_wall"] = e2e * 1e3
            res[mname]["e2e_d2h_bytes"] = int(host_mix.nbytes)
            print(f"{name:34s} end to end (device spans -> fused kernel -> mix D2H, wall clock): {e2e * 1e3:8.3f} ms", flush=True)
        out.free()
        bank.graph.free()
    for b in prepared:
        b.free()
    return res


class NiceBankAdapter:
    """configs[1] through the un-fused NiceInstrumentBank (engine.py), for comparison with the fused render."""

    def __init__(self, ctx, wl, sr, total):
        from zang_b200.events import BankTrigger
        self.ctx, self.voices = ctx, wl["color"].size
        self.bank = E.NiceInstrumentBank(ctx, self.voices, 0.3)
        self.trigger = BankTrigger(wl["event_off"], wl["events"], total, 1024)
        ev = wl["events"]
        self.rows = np.zeros(ev.size, dtype=E.NiceInstrumentBank.Params)
        self.rows["sample_rate"], self.rows["freq"], self.rows["note_on"] = sr, ev["freq"], ev["note_on"]

    def prepare_block(self, first, n):
        off, spans = self.trigger.spans(first, n)
        return (self.ctx.spans(self.voices, n, off, spans),)

    def paint_block(self, first, n, out, temps, prepared=None):
        self.bank.paint(prepared[0], [out], temps, self.rows)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "bank_bench.json"))
    ap.add_argument("--scale", type=float, default=1.0, help="scale the voice counts (smoke runs)")
    ap.add_argument("--only", default="", help="comma list of config numbers to run (1,2,3,4)")
    ap.add_argument("--exact", action="store_true", help="ZCU_PAINT_EXACT (default here: ZCU_PAINT_TOLERANCE, Filter in verified time segments)")
    args = ap.parse_args()
    peak = 6553.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    ctx = E.Context(0)
    ctx.set_paint_mode(L.PAINT_EXACT if args.exact else L.PAINT_TOLERANCE)
    res = []
    sc = args.scale
    only = set(int(x) for x in args.only.split(",")) if args.only else {1, 2, 3, 4}
    # configs[1] polyphony bank, un-fused: 4096 voices, 64 mix buffers
    V, frames = int(4096 * sc), 64 * 1024
    if 1 in only:
        wl = synth.polyphony_bank(V, frames, 48000.0)
        res.append(run(ctx, "configs[1] NiceInstrument (un-fused)", NiceBankAdapter(ctx, wl, 48000.0, frames), frames, 2, 64, peak))
    # configs[2] fm/detuned: 16384 voices, 16 mix buffers
    V, frames = int(16384 * sc), 16 * 1024
    if 2 in only:
        wl = synth.fm_bank(V, frames, 48000.0)
        det, idx = synth.fm_mod_inputs(V, frames)
        ins = (ctx.buf(frames, V, det), ctx.buf(frames, V, idx))
        del det, idx
        res.append(run(ctx, "configs[2] FmDetuned", banks.FmDetunedBank(ctx, wl, 48000.0, frames), frames, 6, 140, peak, ins=ins))
        res[-1].update(run_fused(ctx, "configs[2] FmDetuned", banks.FmDetunedBank(ctx, wl, 48000.0, frames), frames, 140, peak, ins=ins,
                                 gains=np.full((V, 2), 1.0 / V, dtype=np.float32), wl=wl))
        for b in ins:
            b.free()
    # configs[3] sampler: 2048 voices, 64 mix buffers @44.1 kHz
    V, frames = int(2048 * sc), 64 * 1024
    if 3 in only:
        wl = synth.sampler_bank(V, frames, 44100.0)
        pcm, _ = synth.drumloop_pcm()
        smp = ctx.sample(pcm, 1, 44100, L.FMT_SIGNED16_LSB)
        res.append(run(ctx, "configs[3] SamplerVoice", banks.SamplerVoiceBank(ctx, wl, smp, 44100.0, frames), frames, 2, 48, peak))
        res[-1].update(run_fused(ctx, "configs[3] SamplerVoice", banks.SamplerVoiceBank(ctx, wl, smp, 44100.0, frames), frames, 48, peak,
                                 gains=np.full((V, 2), 1.0 / V, dtype=np.float32), wl=wl))
    # configs[4] SFX batch: 65536 jobs x 24 mix buffers
    V, frames = int(65536 * sc), 24 * 1024
    if 4 in only:
        wl = synth.sfx_bank(V, frames)
        res.append(run(ctx, "configs[4] Sfx", banks.SfxBank(ctx, wl, 48000.0, frames), frames, 3, 100, peak))
        res[-1].update(run_fused(ctx, "configs[4] Sfx", banks.SfxBank(ctx, wl, 48000.0, frames), frames, 100, peak, gains=wl["gains"], wl=wl))
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(dict(peak_gbs=peak, paint_mode="exact" if args.exact else "tolerance", results=res), open(args.out, "w"), indent=1)
    ctx.close()


if __name__ == "__main__":
    main()
