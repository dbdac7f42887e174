"""Per-module throughput of the drop-in paint path (one C-ABI call = every span of every voice of a block).

For every module / ConstantOrBuffer variant: a bank of V voices paints NB full 1024-frame spans per voice
(the un-triggered call pattern of examples/example_polyphony.zig:147-160), CUDA-event timed on the context
stream, L2 flushed between repetitions.  Reported: voice-samples/s and the algorithmic HBM GB/s
(SURVEY.md §8a bytes per voice-sample: read-modify-write of the output + one read per input buffer).

  python tools/module_bw.py [--voices 4096,65536] [--buffers 8] [--out gpurun_out/module_bw.json]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from zang_b200 import _lib as L  # noqa: E402
from zang_b200 import engine as E  # noqa: E402

SR = 48000.0


def module_cases():
    """(name, kind, bytes per voice-sample, number of input buffers, params row, kwargs names of the input buffers)"""
    def row(kind, **kw):
        p = np.zeros(1, dtype=L.param_dtypes[kind])
        for k, v in kw.items():
            p[k] = v
        return p
    env = dict(sample_rate=SR, attack=(L.CURVE_CUBED, 0.01), decay=(L.CURVE_CUBED, 0.1), release=(L.CURVE_CUBED, 0.5), sustain_volume=0.8, note_on=1)
    return [
        ("SineOsc const", L.MOD_SINE_OSC, 8, row(L.MOD_SINE_OSC, sample_rate=SR, freq=440.0, phase=0.0), []),
        ("SineOsc freq+phase buf", L.MOD_SINE_OSC, 16, row(L.MOD_SINE_OSC, sample_rate=SR, freq=440.0, phase=0.0), ["freq", "phase"]),
        ("PulseOsc const", L.MOD_PULSE_OSC, 8, row(L.MOD_PULSE_OSC, sample_rate=SR, freq=440.0, color=0.3), []),
        ("PulseOsc freq buf", L.MOD_PULSE_OSC, 12, row(L.MOD_PULSE_OSC, sample_rate=SR, freq=440.0, color=0.3), ["freq"]),
        ("TriSawOsc const", L.MOD_TRISAW_OSC, 8, row(L.MOD_TRISAW_OSC, sample_rate=SR, freq=440.0, color=0.3), []),
        ("TriSawOsc freq buf", L.MOD_TRISAW_OSC, 12, row(L.MOD_TRISAW_OSC, sample_rate=SR, freq=440.0, color=0.3), ["freq"]),
        ("Noise white", L.MOD_NOISE, 8, row(L.MOD_NOISE, color=L.NOISE_WHITE), []),
        ("Noise pink", L.MOD_NOISE, 8, row(L.MOD_NOISE, color=L.NOISE_PINK), []),
        ("Envelope", L.MOD_ENVELOPE, 8, row(L.MOD_ENVELOPE, **env), []),
        ("Gate", L.MOD_GATE, 8, row(L.MOD_GATE, note_on=1), []),
        ("Filter const", L.MOD_FILTER, 12, row(L.MOD_FILTER, type=L.FILTER_LOW_PASS, cutoff=0.2, res=0.7), ["input"]),
        ("Filter cutoff+res buf", L.MOD_FILTER, 20, row(L.MOD_FILTER, type=L.FILTER_LOW_PASS, cutoff=0.2, res=0.7), ["input", "cutoff", "res"]),
        ("Portamento", L.MOD_PORTAMENTO, 8, row(L.MOD_PORTAMENTO, sample_rate=SR, curve=(L.CURVE_LINEAR, 0.5), goal=440.0, note_on=1, prev_note_on=1), []),
        ("Decimator", L.MOD_DECIMATOR, 12, row(L.MOD_DECIMATOR, sample_rate=SR, fake_sample_rate=6000.0), ["input"]),
        ("Distortion overdrive", L.MOD_DISTORTION, 12, row(L.MOD_DISTORTION, type=L.DIST_OVERDRIVE, ingain=0.9, outgain=0.5, offset=0.0), ["input"]),
        ("Distortion clip", L.MOD_DISTORTION, 12, row(L.MOD_DISTORTION, type=L.DIST_CLIP, ingain=0.9, outgain=0.5, offset=0.0), ["input"]),
        ("Cycle const", L.MOD_CYCLE, 8, row(L.MOD_CYCLE, sample_rate=SR, speed=3.0), []),
    ]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--voices", default="4096,65536")
    ap.add_argument("--buffers", type=int, default=8)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--only", default="", help="comma-separated substring filters on the module name")
    ap.add_argument("--tolerance", action="store_true", help="ZCU_PAINT_TOLERANCE: Filter cut into verified time segments")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "module_bw.json"))
    args = ap.parse_args()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6553.0))
    ctx = E.Context(0)
    if args.tolerance:
        ctx.set_paint_mode(L.PAINT_TOLERANCE)
    results = []
    for V in [int(x) for x in args.voices.split(",")]:
        NB = args.buffers
        frames = NB * 1024
        rng = np.random.default_rng(1)
        sp = ctx.spans_full(V, NB, 1024)
        out = ctx.buf(frames, V)
        flush = ctx.buf(64 * 1024, 1024)  # 256 MiB > L2
        ins = {}
        host = rng.uniform(0.05, 0.6, size=(frames, V)).astype(np.float32)
        for name in ("freq", "phase", "input", "cutoff", "res"):
            ins[name] = ctx.buf(frames, V)
            scale = 2000.0 if name == "freq" else 1.0
            ins[name].upload(host * np.float32(scale))
        del host
        smp = None
        for name, kind, bps, prow, inputs in module_cases():
            if args.only and not any(o in name for o in args.only.split(",")):
                continue
            mod = E.MODULE_CLASSES[kind](ctx, V)
            kw = {k: ins[k] for k in inputs}
            ms_best = 1e30
            for rep in range(args.reps + 1):
                ctx.zero(None, flush)
                ctx.zero(None, out)
                ctx.timer_start()
                mod.paint(sp, [out], [], prow, **kw)
                ms = ctx.timer_stop_ms()
                if rep > 0:
                    ms_best = min(ms_best, ms)
            vs = V * frames / (ms_best * 1e-3)
            gbs = vs * bps / 1e9
            results.append(dict(module=name, voices=V, frames=frames, ms=ms_best, voice_samples_per_s=vs, bytes_per_voice_sample=bps, gbs=gbs, frac_of_peak=gbs / peak))
            print(f"V={V:6d} {name:26s} {ms_best:9.3f} ms  {vs:10.3e} vs/s  {gbs:8.1f} GB/s  {gbs / peak:6.3f} of {peak:.0f}", flush=True)
        # modules with extra arguments (nodes, PCM table, delay lines): same timing, explicit calls
        from zang_b200 import synth
        pcm = ctx.sample(synth.pcm_loop(), 1, 44100, L.FMT_SIGNED16_LSB)
        nodes = np.array([(1000.0, 0.0), (200.0, 0.1), (100.0, 0.2), (400.0, 5.0)], dtype=L.curve_node_dtype)
        curve = E.Curve(ctx, V)
        smp = E.Sampler(ctx, V)
        sdel = E.SimpleDelay(ctx, V, 7500)
        fecho = E.FilteredEchoes(ctx, V, 15000)
        sdel2 = E.SimpleDelay(ctx, V, 15000) if "SimpleDelay 15000" in args.only else None
        fmo = E.FmOscillator(ctx, V)

        def prow(kind, **kw):
            p = np.zeros(1, dtype=L.param_dtypes[kind])
            for k, v in kw.items():
                p[k] = v
            return p
        custom = [
            ("Curve smoothstep", 8, lambda: curve.paint(sp, [out], [], prow(L.MOD_CURVE, sample_rate=SR, function=L.INTERP_SMOOTHSTEP, node_start=0, node_count=4), nodes)),
            ("Sampler resample x0.73", 8, lambda: smp.paint(sp, [out], [], prow(L.MOD_SAMPLER, sample_rate=44100.0 / 0.73, channel=0, loop=1), pcm)),
            ("SimpleDelay 7500", 20, lambda: sdel.paint(sp, [out], [], ins["input"])),
            ("SimpleDelay 15000", 20, (lambda: sdel2.paint(sp, [out], [], ins["input"])) if sdel2 else None),
            ("FilteredEchoes 15000", 20, lambda: fecho.paint(sp, [out], [], prow(L.MOD_FILTERED_ECHOES, feedback_volume=0.6, cutoff=0.1), ins["input"])),
            ("FmOscillator freq buf", 12, lambda: fmo.paint(sp, [out], [], prow(L.MOD_FM_OSCILLATOR, sample_rate=SR, waveform=0, feedback=0.2), ins["freq"], None)),
        ]
        for name, bps, fn in custom:
            if fn is None or (args.only and not any(o in name for o in args.only.split(","))):
                continue
            ms_best = 1e30
            for rep in range(args.reps + 1):
                ctx.zero(None, flush)
                ctx.zero(None, out)
                ctx.timer_start()
                fn()
                ms = ctx.timer_stop_ms()
                if rep > 0:
                    ms_best = min(ms_best, ms)
            vs = V * frames / (ms_best * 1e-3)
            gbs = vs * bps / 1e9
            results.append(dict(module=name, voices=V, frames=frames, ms=ms_best, voice_samples_per_s=vs, bytes_per_voice_sample=bps, gbs=gbs, frac_of_peak=gbs / peak))
            print(f"V={V:6d} {name:26s} {ms_best:9.3f} ms  {vs:10.3e} vs/s  {gbs:8.1f} GB/s  {gbs / peak:6.3f} of {peak:.0f}", flush=True)
        # whole-buffer basics (spans == NULL)
        a, b = ins["input"], ins["cutoff"]
        for name, bps, fn in [] if args.only else [("zero", 4, lambda: ctx.zero(None, out)), ("copy", 8, lambda: ctx.copy(None, out, a)),
                              ("addInto", 12, lambda: ctx.addInto(None, out, a)), ("multiply", 16, lambda: ctx.multiply(None, out, a, b)),
                              ("multiplyWithScalar", 8, lambda: ctx.multiplyWithScalar(None, out, 0.5))]:
            ms_best = 1e30
            for rep in range(args.reps + 1):
                ctx.zero(None, flush)
                ctx.timer_start()
                fn()
                ms = ctx.timer_stop_ms()
                if rep > 0:
                    ms_best = min(ms_best, ms)
            vs = V * frames / (ms_best * 1e-3)
            gbs = vs * bps / 1e9
            results.append(dict(module="basics." + name, voices=V, frames=frames, ms=ms_best, voice_samples_per_s=vs, bytes_per_voice_sample=bps, gbs=gbs, frac_of_peak=gbs / peak))
            print(f"V={V:6d} {'basics.' + name:26s} {ms_best:9.3f} ms  {vs:10.3e} vs/s  {gbs:8.1f} GB/s  {gbs / peak:6.3f} of {peak:.0f}", flush=True)
        for x in list(ins.values()) + [out, flush, sp]:
            x.free()
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(dict(peak_gbs=peak, paint_mode="tolerance" if args.tolerance else "exact", results=results), open(args.out, "w"), indent=1)
    ctx.close()


if __name__ == "__main__":
    main()
