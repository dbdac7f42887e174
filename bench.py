#!/usr/bin/env python
"""bench.py — voice-samples/s of the polyphony-bank render (BASELINE.json configs[1]) on N B200s.

Workload ("step" = one whole render job): 4096 voices per GPU of NiceInstrument
(PulseOsc -> x0.5 -> Filter low-pass res 0.7 -> x Envelope; examples/modules.zig:189-248), each voice
driven by its own synthetic note stream through Trigger span splitting, 60 s at 48 kHz in 1024-frame
mix buffers, summed to a stereo mix with per-voice pan gains.  With N > 1 each rank renders its own
4096 voices (weak scaling: voices are independent, no data-path exchange) and the per-GPU partial
mixes are summed with ONE NCCL reduce to rank 0 inside the timed region.

  value  : voice-samples/s, device timed (CUDA events on the launch stream), inputs resident in HBM
  e2e    : same metric through the C ABI from HOST buffers: H2D of the note events (pinned), render,
           (reduce,) D2H of the stereo mix (pinned) — all inside the timed region, every step; jobs are
           submitted back to back, so a job's upload and the previous job's D2H overlap the render
  roofline: algorithmic bytes (64 B per voice-sample, SURVEY.md §8d: what the reference's module chain
           moves through its buffers) of the dominant kernel / its event-timed duration, vs the measured
           HBM copy peak.  The fused kernel keeps the chain in registers, so `traffic` (ncu DRAM bytes)
           is far below the algorithmic bytes and frac may exceed 1; DESIGN.md explains.
  cpu_baseline: the oracle (C++ restatement of the reference loops; no Zig toolchain exists here) on the
           host cores, on a bounded sample of the same workload.  Reported, not the target.

  parity : the benched job checked, untimed, against the oracle on rank 0 (tests/bench_parity.py: per-voice windows at
           every segment start + random windows + the first 64 voices x 8 s + the stereo mix)
  strong : the same 4096-voice job split V/N per GPU (strong scaling), one reduce of the mix
  sfx    : BASELINE.json configs[4]: 65536 SFX jobs x 24576 frames split over the GPUs, one reduce of the mix
  other_configs : configs[2], configs[3] through the drop-in module path (N = 1 only), for a driver clock record

`--impl reference` times that CPU path alone (rank 0 only).
"""
import argparse
import json
import os

# The driver reads ONE JSON line from stdout.  NCCL prints its version banner on stdout (NCCL_DEBUG=VERSION and WARN both do),
# and other libraries may chat too: file descriptor 1 is pointed at stderr for the whole run and the result line is written
# to the saved descriptor by emit().
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(line):
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())

import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

VOICES_PER_GPU = 4096
SECONDS = 60
SAMPLE_RATE = 48000.0
BUFFER_FRAMES = 1024
ALGO_BYTES_PER_VOICE_SAMPLE = 64  # zero4+Pulse8+mulScalar8+zero4+Filter12+zero4+Env8+multiply16
METRIC = "voice-samples/sec"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--voices", type=int, default=VOICES_PER_GPU, help="voices per GPU (default: the BASELINE config)")
    ap.add_argument("--seconds", type=float, default=SECONDS)
    ap.add_argument("--block-buffers", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-other-modes", action="store_true", help="skip the comparison runs of the other render modes")
    ap.add_argument("--no-parity", action="store_true", help="skip the untimed oracle check of the benched job")
    ap.add_argument("--no-extra", action="store_true", help="skip the strong-scaling, SFX and other-config legs")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak (default, the driver's line): 4096 voices per GPU; strong: 4096 voices in total, V/N per GPU")
    ap.add_argument("--config", default="polyphony", choices=["polyphony", "sfx"],
                    help="sfx: BASELINE.json configs[4] as the headline line (65536 jobs x 24576 frames split over the GPUs)")
    ap.add_argument("--mode", default="segments", choices=["exact", "scan", "segments"],
                    help="exact: sequential-in-time kernel, bit-exact per voice; scan: time-parallel chunks (exact PulseOsc/Envelope, "
                         "filter start states from a 2x2 f64 scan; per-voice error <= 1e-5, the north-star tolerance)")
    return ap.parse_args()


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json, torch copy)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic_bytes_per_launch():
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture, if there is one."""
    p = os.path.join(ROOT, "profiles", "dominant_kernel_traffic.json")
    try:
        with open(p) as f:
            return json.load(f)
    except Exception:
        return None


MODE_INFO = {
    "exact": ("k_render_nice", "k_render_nice (fused PulseOsc+Filter+Envelope+mix, Trigger replay on device, sequential in time)",
              "exact: every voice sequential in time; per-voice output bit-exact vs the oracle"),
    "scan": ("k_nice_scan_render", "k_nice_scan_render (pass 3 of zero-state pass / 2x2 f64 chain / render; one warp per 32 voices x mix buffer)",
             "scan: one warp task per (32 voices, mix buffer); PulseOsc/Envelope exact closed forms, Filter start states from a chained "
             "2x2 f64 scan; per-voice error <= 1e-5 of full scale (tests/test_gpu_instrument.py)"),
    "segments": ("k_nice_seg_render", "k_nice_seg_render (fused PulseOsc+Filter+Envelope+mix; (32 voices x time segment) per warp, Filter warm-up "
                 "verified by the window's homogeneous map)",
                 "segments: time axis cut into GPU-filling segments rendered sequentially; PulseOsc/Envelope exact closed forms, Filter state "
                 "by warm-up until the window map <= 1e-8; per-voice error <= 1e-5 of full scale (tests/test_gpu_instrument.py)"),
}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except Exception:
                continue
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            for name, val in zip(names, r[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_reference_run(steps, warmup, quiet=False):
    """The reference's CPU path (oracle port) on all host threads, bounded sample per step."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle  # noqa: E402  (bench's cpu_baseline / reference legs are allowed to execute oracle/)
    from zang_b200 import synth
    threads = max(1, min(32, len(os.sched_getaffinity(0))))
    voices = 32 * threads
    total = int(SECONDS * SAMPLE_RATE)
    wl = synth.polyphony_bank(voices, total, SAMPLE_RATE)
    times = []
    for i in range(warmup + steps):
        _, _, secs = oracle.render_nice(wl, total, SAMPLE_RATE, BUFFER_FRAMES, channels=2, nthreads=threads)
        if i >= warmup:
            times.append(secs)
    t = sum(times) / len(times)
    value = voices * total / t
    sample = f"{voices} voices x {SECONDS} s @48 kHz NiceInstrument bank (same generator as the GPU workload), {threads} threads"
    return {"value": value, "unit": METRIC, "cores": threads, "kind": "port", "sample": sample, "ms_per_step": t * 1e3,
            "note": "C++ restatement of the reference's Zig paint loops (oracle/); the Zig ReleaseFast build cannot be produced here"}


def sfx_leg(ctx, torch, dist, timed, rank, world, peak, jobs_total=65536, job_frames=24 * 1024):
    """BASELINE.json configs[4]: jobs_total SFX jobs (Noise -> Curve -> Filter(cutoff buffer) -> Envelope -> multiply;
    example_stereo.zig:71-98, example_laser.zig:77-115) split jobs/N per GPU, per-job pan gains, one reduce of the stereo
    mix.  Device timed: module chain + voice sum + reduce; span tables and params resident."""
    from zang_b200 import banks, synth
    wl_all = synth.sfx_bank(jobs_total, job_frames)
    a, b = jobs_total * rank // world, jobs_total * (rank + 1) // world
    wl = {"event_off": (wl_all["event_off"][a:b + 1] - wl_all["event_off"][a]).astype("uint32"),
          "events": wl_all["events"][wl_all["event_off"][a]:wl_all["event_off"][b]], "filter_type": wl_all["filter_type"][a:b],
          "nodes": wl_all["nodes"].reshape(jobs_total, 3)[a:b].reshape(-1), "gains": wl_all["gains"][a:b]}
    V = b - a
    bank = banks.SfxBank(ctx, wl, SAMPLE_RATE, job_frames, first_seed=a)
    out = ctx.buf(job_frames, V)
    temps = [ctx.buf(job_frames, V) for _ in range(3)]
    mix_t = torch.zeros((job_frames, 2), dtype=torch.float32, device="cuda")
    mix = ctx.wrap(mix_t.data_ptr(), job_frames, 2)
    prepared = bank.prepare_block(0, job_frames)
    from zang_b200 import _lib as L
    ctx.set_paint_mode(L.PAINT_TOLERANCE)

    def step():
        bank.noise.bank.reset(a)  # every step renders the same jobs from init()
        for m in (bank.curve, bank.filter, bank.env):
            m.bank.reset()
        ctx.zero(None, out)
        bank.paint_block(0, job_frames, out, temps, prepared=prepared)
        mix_t.zero_()
        ctx.mix_reduce(out, wl["gains"], 2, mix)
        if dist is not None:
            dist.reduce(mix_t, dst=0)
    l0 = ctx.launch_count
    step()
    launches = ctx.launch_count - l0
    ms = timed(step, 3)
    ctx.set_paint_mode(L.PAINT_EXACT)
    dropin_mix = mix_t.clone() if rank == 0 else None

    # the same jobs as ONE fused graph kernel (zcu_graph_*, MIX output: Noise -> Curve -> Filter -> x Envelope -> pan + voice sum
    # with the temps in registers; bit-identical per job to the module chain, tests/test_gpu_graph.py)
    bank.build_graph(output_mode=L.GRAPH_OUT_MIX, channels=2, gains=wl["gains"])

    def step_fused():
        bank.noise.bank.reset(a)
        for m in (bank.curve, bank.filter, bank.env):
            m.bank.reset()
        mix_t.zero_()
        bank.paint_block_fused(0, job_frames, mix, prepared=prepared)
        if dist is not None:
            dist.reduce(mix_t, dst=0)
    # tolerance mode, as for the module chain above: with fewer jobs per GPU than fill it, the graph is cut into time segments
    # (the Filter starts each from a verified warm-up window, tests/test_gpu_graph.py::test_sfx_graph_filter_time_segments)
    ctx.set_paint_mode(L.PAINT_TOLERANCE)
    l0 = ctx.launch_count
    step_fused()
    launches_f = ctx.launch_count - l0
    ms_f = timed(step_fused, 3)
    seg_info = bank.graph.segment_info()
    seg_info["warm_retries"] = bank.graph.warm_retries()
    ctx.set_paint_mode(L.PAINT_EXACT)
    fused_vs_dropin = float((mix_t - dropin_mix).abs().max()) if rank == 0 else 0.0
    mix_scale = float(dropin_mix.abs().max()) if rank == 0 else 0.0
    t = torch.tensor([sum(ms) / len(ms), sum(ms_f) / len(ms_f)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step, ms_fused = float(t[0]), float(t[1])
    bank.graph.free()
    for x in [out] + temps + list(prepared):
        x.free()
    vs = float(jobs_total) * job_frames / (ms_fused * 1e-3)
    vs_d = float(jobs_total) * job_frames / (ms_step * 1e-3)
    gbs = vs_d * 100 / 1e9
    return {"workload": "BASELINE.json configs[4]: SFX jobs (Noise -> Curve -> Filter with cutoff buffer -> x Envelope), pan + voice sum, one NCCL "
                        "reduce of the mix; value = the fused graph kernel (zcu_graph_*), dropin = the module chain (ZCU_PAINT_TOLERANCE)",
            "jobs_total": jobs_total, "jobs_per_gpu": V, "job_frames": job_frames, "n_gpus": world, "scaling": "strong",
            "ms_per_step": ms_fused, "value": vs, "unit": METRIC, "steps": 3, "gpu_launches_per_step": int(launches_f),
            "fused_time_segments": seg_info,
            "mix_max_abs_diff_fused_vs_dropin": fused_vs_dropin, "mix_abs_max": mix_scale,
            "dropin": {"ms_per_step": ms_step, "value": vs_d, "gpu_launches_per_step": int(launches), "algorithmic_bytes_per_voice_sample": 100,
                       "algorithmic_gbs_all_gpus": gbs, "frac_of_hbm_peak_per_gpu": gbs / world / peak}}


def other_configs_leg(ctx, torch, timed, peak):
    """configs[2] (FmDetuned, 16384 voices) and configs[3] (SamplerVoice, 2048 voices): one block each through the drop-in
    module path (device timed, span tables resident) and as one fused graph kernel with the span tables built on the
    device (N = 1 only).  tools/bank_bench.py is the long form."""
    import numpy as np
    from zang_b200 import _lib as L
    from zang_b200 import banks, synth
    res = {}
    ctx.set_paint_mode(L.PAINT_TOLERANCE)

    def fused(bank, wl, V, frames, ins=()):
        """The same body as ONE fused graph kernel (zcu_graph_*), span tables built on the device (zcu_event_bank_*)."""
        bank.use_device_spans(wl)
        for x in bank.prepare_block(0, frames):  # warm-up: allocations, static tables
            x.free()
        bank.dev_trigger.reset()
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        torch.cuda.synchronize()
        ev[0].record(torch.cuda.current_stream())
        prep = bank.prepare_block(0, frames)
        ev[1].record(torch.cuda.current_stream())
        torch.cuda.synchronize()
        t_build = time.perf_counter()
        bank.build_graph()
        t_build = time.perf_counter() - t_build
        fout = ctx.buf(frames, V)

        def step_f():
            bank.paint_block_fused(0, frames, fout, *ins, prepared=prep)
        step_f()
        msf = timed(step_f, 3)
        mf = sum(msf) / len(msf)
        bank.graph.free()
        for x in [fout] + list(prep):
            x.free()
        return {"fused_graph_ms_per_block": mf, "fused_graph_value": V * frames / (mf * 1e-3), "graph_build_s": t_build,
                "device_span_tables_ms": ev[0].elapsed_time(ev[1])}
    try:
        # configs[2]
        V, frames = 16384, 16 * 1024
        wl = synth.fm_bank(V, frames, SAMPLE_RATE)
        f = torch.arange(frames, dtype=torch.float64, device="cuda")[:, None]
        v = torch.arange(V, dtype=torch.float64, device="cuda")[None, :]
        det = (1.0 + 0.02 * torch.sin(2 * np.pi * (0.37 + 0.013 * (v % 17)) * f / 48000.0 + 0.7 * v)).float().contiguous()
        idx = (1.0 + torch.sin(2 * np.pi * (0.11 + 0.007 * (v % 23)) * f / 48000.0 + 1.3 * v)).float().contiguous()
        ins = (ctx.wrap(det.data_ptr(), frames, V), ctx.wrap(idx.data_ptr(), frames, V))
        bank = banks.FmDetunedBank(ctx, wl, SAMPLE_RATE, frames)
        out = ctx.buf(frames, V)
        temps = [ctx.buf(frames, V) for _ in range(6)]
        prepared = bank.prepare_block(0, frames)

        def step_fm():
            bank.paint_block(0, frames, out, temps, *ins, prepared=prepared)
        step_fm()
        ms = timed(step_fm, 3)
        m = sum(ms) / len(ms)
        gbs = V * frames * 140 / (m * 1e-3) / 1e9
        res["configs[2] FmDetuned 16384 voices x 16384 frames"] = {"ms_per_block": m, "value": V * frames / (m * 1e-3), "bytes_per_voice_sample": 140,
                                                                  "algorithmic_gbs": gbs, "frac_of_hbm_peak": gbs / peak}
        for x in [out] + temps + list(prepared):
            x.free()
        res["configs[2] FmDetuned 16384 voices x 16384 frames"].update(fused(banks.FmDetunedBank(ctx, wl, SAMPLE_RATE, frames), wl, V, frames, ins))
        del det, idx
        # configs[3]
        V, frames = 2048, 64 * 1024
        wl = synth.sampler_bank(V, frames, 44100.0)
        pcm, pcm_name = synth.drumloop_pcm()
        smp = ctx.sample(pcm, 1, 44100, L.FMT_SIGNED16_LSB)
        bank = banks.SamplerVoiceBank(ctx, wl, smp, 44100.0, frames)
        out = ctx.buf(frames, V)
        temps = [ctx.buf(frames, V) for _ in range(2)]
        prepared = bank.prepare_block(0, frames)

        def step_smp():
            bank.paint_block(0, frames, out, temps, prepared=prepared)
        step_smp()
        ms = timed(step_smp, 3)
        m = sum(ms) / len(ms)
        gbs = V * frames * 48 / (m * 1e-3) / 1e9
        res[f"configs[3] SamplerVoice 2048 voices x 65536 frames ({pcm_name})"] = {"ms_per_block": m, "value": V * frames / (m * 1e-3),
                                                                                 "bytes_per_voice_sample": 48, "algorithmic_gbs": gbs,
                                                                                 "frac_of_hbm_peak": gbs / peak}
        for x in [out] + temps + list(prepared):
            x.free()
        res[f"configs[3] SamplerVoice 2048 voices x 65536 frames ({pcm_name})"].update(
            fused(banks.SamplerVoiceBank(ctx, wl, smp, 44100.0, frames), wl, V, frames))
    except Exception as e:  # never at the expense of the headline
        res["error"] = repr(e)
    ctx.set_paint_mode(L.PAINT_EXACT)
    return res


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    total_frames = int(args.seconds * SAMPLE_RATE)

    if args.impl == "reference":
        if rank != 0:
            return 0
        res = cpu_reference_run(max(1, args.steps), max(0, args.warmup))
        line = {"impl": "reference", "metric": METRIC, "value": res["value"], "unit": METRIC, "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "config": {"workload": "polyphony bank: NiceInstrument (PulseOsc+Envelope+Filter lowpass res 0.7), 60 s @48 kHz, stereo mix",
                           "voices_per_gpu": VOICES_PER_GPU, "note": "CPU reference arm renders a bounded sample of the same workload"},
                "cpu_baseline": {k: res[k] for k in ("value", "unit", "cores", "kind", "sample")},
                "e2e": {"value": res["value"], "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        emit(line)
        return 0

    import numpy as np
    import torch
    from zang_b200 import _lib as L
    from zang_b200 import synth
    from zang_b200.engine import Context, Render

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libzangcu has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist = dist_mod
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    stream = torch.cuda.current_stream()
    ctx = Context(local_rank, stream=stream.cuda_stream)  # every kernel, copy, event and the NCCL reduce share one stream
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    peak, peak_src = measured_peak()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        """K steps between barrier+synchronize; per-step device time from CUDA events on the launch stream;
        L2 flushed (untimed) between steps."""
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        barrier()
        for a, b in evs:
            flush.zero_()
            a.record(stream)
            fn()
            b.record(stream)
        barrier()
        return [a.elapsed_time(b) for a, b in evs]

    if args.config == "sfx":  # configs[4] as the line itself
        sfx = sfx_leg(ctx, torch, dist, timed, rank, world, peak)
        if rank == 0:
            emit({"metric": METRIC, "value": sfx["value"], "unit": METRIC, "n_gpus": world, "steps": 3, "warmup": 1, "ms_per_step": sfx["ms_per_step"],
                  "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": sfx,
                  "gpu_launches": sfx["gpu_launches_per_step"] * 3})
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    # ---- workload: weak = the global bank has voices*world voices, strong = `voices` in total; a rank owns a contiguous shard
    strong_line = args.scaling == "strong"
    V_global = args.voices if strong_line else args.voices * world
    wl_all = synth.polyphony_bank(V_global, total_frames, SAMPLE_RATE)
    wl = synth.shard_voices(wl_all, rank, world)
    V = int(wl["color"].size)
    n_events = int(wl["event_off"][-1])
    # pinned host buffers for the end-to-end leg
    ev_pinned = torch.empty(max(n_events, 1) * L.note_event_dtype.itemsize, dtype=torch.uint8).pin_memory()
    ev_host = ev_pinned.numpy().view(L.note_event_dtype)[:n_events]
    ev_host[:] = wl["events"]
    off_pinned = torch.empty(V + 1, dtype=torch.int32).pin_memory()
    off_host = off_pinned.numpy().view(np.uint32)
    off_host[:] = wl["event_off"]
    mix_host = torch.empty((total_frames, 2), dtype=torch.float32).pin_memory()

    mix_t = torch.zeros((total_frames, 2), dtype=torch.float32, device="cuda")
    mix = ctx.wrap(mix_t.data_ptr(), total_frames, 2)
    mode_id = {"exact": L.RENDER_EXACT, "scan": L.RENDER_SCAN, "segments": L.RENDER_SEGMENTS}[args.mode]
    r = Render(ctx, V, total_frames, wl["color"], wl["gains"], channels=2, sample_rate=SAMPLE_RATE, buffer_frames=BUFFER_FRAMES,
               block_buffers=args.block_buffers, mode=mode_id)
    r.set_events(off_host, ev_host)

    def step_resident():
        mix_t.zero_()
        r.run(mix)
        if dist is not None:
            dist.reduce(mix_t, dst=0)

    # end-to-end leg: jobs are submitted back to back, so a job's inputs (set_events: host expansion, H2D and pre-pass on the
    # render's upload stream) and the previous job's result (D2H on a copy stream, two mix buffers) overlap the render.
    # Every step still copies its events in and its mix out; a step's timed interval closes only after the PREVIOUS step's
    # D2H has completed, so in steady state each interval accounts for one whole D2H.
    copy_stream = torch.cuda.Stream()
    mix2_t = [mix_t, torch.zeros_like(mix_t)]
    mix2 = [mix, ctx.wrap(mix2_t[1].data_ptr(), total_frames, 2)]
    mix2_host = [mix_host, torch.empty((total_frames, 2), dtype=torch.float32).pin_memory()]
    e2e_state = {"i": 0, "copy_done": None}

    def step_e2e():
        i = e2e_state["i"] & 1
        e2e_state["i"] += 1
        r.set_events(off_host, ev_host, sync=False)  # H2D from pinned memory (+ host scalars per event)
        mix2_t[i].zero_()
        r.run(mix2[i])
        if dist is not None:
            dist.reduce(mix2_t[i], dst=0)
        prev = e2e_state["copy_done"]
        if rank == 0:
            ready = torch.cuda.Event()
            ready.record(stream)
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(ready)
                mix2_host[i].copy_(mix2_t[i], non_blocking=True)  # D2H of the step's result
                done = torch.cuda.Event()
                done.record(copy_stream)
            e2e_state["copy_done"] = done
        if prev is not None:
            stream.wait_event(prev)  # the other mix buffer is free again, and the interval includes that copy

    for _ in range(max(args.warmup, 0)):
        step_resident()
    barrier()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = ctx.launch_count
    r.set_profiling(True)
    ms_steps = timed(step_resident, args.steps)
    kern_ms, kern_launches = r.kernel_ms()
    r.set_profiling(False)
    launches = ctx.launch_count - launches0
    clocks = sampler.stop() if rank == 0 else None

    for _ in range(min(2, max(args.warmup, 0))):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps)

    # wall-clock cross-checks of the end-to-end number (host clock, not stream events):
    #   pipelined: K back-to-back jobs between two synchronizes, no flush inside (the flush is what the event-timed leg leaves out);
    #   latency:   one job alone, set_events -> run -> (reduce) -> D2H -> synchronize, nothing to overlap with.
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    wall_pipelined_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    lat = []
    for _ in range(3):
        barrier()
        flush.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r.set_events(off_host, ev_host, sync=False)
        mix_t.zero_()
        r.run(mix)
        if dist is not None:
            dist.reduce(mix_t, dst=0)
        if rank == 0:
            mix_host.copy_(mix_t, non_blocking=True)
        torch.cuda.synchronize()
        lat.append((time.perf_counter() - t0) * 1e3)
    e2e_state["copy_done"] = None

    # max over ranks of the mean step time
    t_res = torch.tensor([sum(ms_steps) / len(ms_steps), sum(ms_e2e) / len(ms_e2e), wall_pipelined_ms, min(lat)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t_res, op=dist.ReduceOp.MAX)
    ms_per_step, ms_e2e_step, wall_pipelined_ms, wall_latency_ms = (float(x) for x in t_res)

    # ---- parity of the benched job, untimed: a clean run, the reduced mix read back on rank 0, then the oracle ----
    parity = None
    if not args.no_parity and args.mode == "segments":
        mix_t.zero_()
        r.run(mix)
        if dist is not None:
            dist.reduce(mix_t, dst=0)
        torch.cuda.synchronize()
        if rank == 0:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            try:
                import bench_parity  # executes oracle/ as the checker (allowed: never inside a timed region)
                nthreads = max(1, min(32, len(os.sched_getaffinity(0))))
                got = mix_t.cpu().numpy()
                # with N ranks the mix holds N x 4096 voices: the oracle then renders the first 1/N of the job of all of them
                parity = bench_parity.check_bench_job(ctx, r, wl, total_frames, SAMPLE_RATE, BUFFER_FRAMES, mix_got=got, mix_wl=wl_all,
                                                      mix_frames=total_frames // world, nthreads=nthreads)
            except AssertionError as e:
                parity = {"ok": False, "error": str(e)}
        barrier()

    # ---- strong scaling: the SAME 4096-voice job split V/N per GPU (at N = 1 it is the main line) ----
    strong = None
    if not args.no_extra and not strong_line and args.mode == "segments":
        if world == 1:
            strong = {"voices_total": V, "voices_per_gpu": V, "n_gpus": 1, "ms_per_step": ms_per_step, "value": float(V) * total_frames / (ms_per_step * 1e-3),
                      "unit": METRIC, "scaling": "strong", "note": "N = 1: identical to the main line"}
        else:
            wl_s = synth.shard_voices(synth.polyphony_bank(args.voices, total_frames, SAMPLE_RATE), rank, world)
            Vs = int(wl_s["color"].size)
            rs = Render(ctx, Vs, total_frames, wl_s["color"], wl_s["gains"], channels=2, sample_rate=SAMPLE_RATE, buffer_frames=BUFFER_FRAMES, mode=mode_id)
            rs.set_events(wl_s["event_off"], wl_s["events"])

            def step_strong():
                mix_t.zero_()
                rs.run(mix)
                if dist is not None:
                    dist.reduce(mix_t, dst=0)
            for _ in range(3):
                step_strong()
            rs.set_profiling(True)
            ms_s = timed(step_strong, args.steps)
            ks_ms, ks_n = rs.kernel_ms()
            rs.set_profiling(False)
            seg_s = rs.segment_info()

            def step_strong_e2e():
                rs.set_events(wl_s["event_off"], wl_s["events"], sync=False)
                step_strong()
                if rank == 0:
                    mix_host.copy_(mix_t, non_blocking=True)
            step_strong_e2e()
            ms_se = timed(step_strong_e2e, args.steps)
            ts = torch.tensor([sum(ms_s) / len(ms_s), sum(ms_se) / len(ms_se), ks_ms / max(ks_n, 1)], dtype=torch.float64, device="cuda")
            dist.all_reduce(ts, op=dist.ReduceOp.MAX)
            strong = {"voices_total": args.voices, "voices_per_gpu": Vs, "n_gpus": world, "ms_per_step": float(ts[0]),
                      "value": float(args.voices) * total_frames / (float(ts[0]) * 1e-3), "unit": METRIC, "scaling": "strong", "steps": args.steps,
                      "render_kernel_ms": float(ts[2]), "segments": seg_s,
                      "e2e_ms_per_step": float(ts[1]), "e2e_value": float(args.voices) * total_frames / (float(ts[1]) * 1e-3),
                      "note": "fixed costs per job that do not shrink with N: mix memset + k_partial_reduce over the 23 MB mix, the NCCL reduce of "
                              "23 MB, launch latencies, and the Filter warm-up (2 buffers per segment, segments get shorter as N grows)"}
            rs.free()

    sfx = None
    if not args.no_extra and args.mode == "segments":
        try:
            sfx = sfx_leg(ctx, torch, dist, timed, rank, world, peak)
        except Exception as e:
            sfx = {"error": repr(e)}

    if rank == 0:
        voice_samples = float(V_global) * total_frames
        value = voice_samples / (ms_per_step * 1e-3)
        e2e_value = voice_samples / (ms_e2e_step * 1e-3)
        kernel_key, kernel_desc, mode_desc = MODE_INFO[args.mode]
        avg_launch_ms = kern_ms / max(kern_launches, 1)
        frames_per_launch = total_frames * args.steps / max(kern_launches, 1)
        block_frames = int(round(frames_per_launch))
        algo_bytes_per_launch = ALGO_BYTES_PER_VOICE_SAMPLE * V * frames_per_launch
        achieved = algo_bytes_per_launch / (avg_launch_ms * 1e-3) / 1e9
        traffic = ncu_traffic_bytes_per_launch()
        if traffic and traffic.get("kernel") != kernel_key:
            traffic = None
        # Roofline of the dominant kernel.  The fused kernel keeps the reference's module chain in registers, so its HBM side
        # is the MEASURED DRAM traffic (ncu, committed capture) over the event-timed launch: a true fraction of the copy peak,
        # about 2 %.  What bounds it is instruction issue (`issue`).  The contract's figure - the algorithmic bytes of the
        # reference's UN-FUSED chain (64 B per voice-sample) over the same time - is kept beside it as `algorithmic`: it is
        # what the fusion saves, not a bandwidth, and exceeds the peak.  The HBM-bound form of the workload is
        # dropin_module_path, whose frac_of_hbm_peak is a roofline fraction in the usual sense.
        dram_gbs = (traffic["dram_bytes_per_launch"] / (avg_launch_ms * 1e-3) / 1e9) if traffic else None
        roofline = {"bound": "hbm", "kernel": kernel_desc,
                    "achieved": dram_gbs, "peak": peak, "unit": "GB/s", "frac": (dram_gbs / peak) if dram_gbs is not None else None,
                    "traffic": traffic["dram_bytes_per_launch"] if traffic else None,
                    "traffic_source": traffic.get("source") if traffic else None,
                    "avg_launch_ms": avg_launch_ms, "launches_timed": int(kern_launches), "kernel_share_of_step": kern_ms / sum(ms_steps),
                    "peak_source": peak_src,
                    "limiter": "instruction issue (the kernel is compute-side bound; HBM is 2 % busy)",
                    "issue": traffic.get("issue") if traffic else None,
                    "algorithmic": {"bytes_per_voice_sample": ALGO_BYTES_PER_VOICE_SAMPLE, "bytes_per_launch": algo_bytes_per_launch,
                                    "gbs": achieved, "frac_of_peak": achieved / peak,
                                    "note": "SURVEY.md 8(d): bytes the reference's un-fused module chain moves through its buffers, over the fused "
                                            "kernel's time; > 1 is the fusion saving, not a bandwidth"}}
        h2d = n_events * 32 + (V + 1) * 4  # device event records (note event + host-evaluated cutoff, 32 B) + offsets
        d2h = total_frames * 2 * 4
        line = {
            "metric": METRIC, "value": value, "unit": METRIC, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if strong_line else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": "polyphony bank: NiceInstrument (PulseOsc+Envelope+Filter lowpass res 0.7) x Trigger note streams, "
                                   f"{args.seconds:g} s @48 kHz in 1024-frame buffers, stereo mix (BASELINE.json configs[1])",
                       "voices_per_gpu": V, "voices_total": V_global, "frames": total_frames, "block_frames": block_frames,
                       "parallelism": f"voices sharded over {world} GPU(s); one NCCL reduce of the stereo mix per job" if world > 1 else "1 GPU",
                       "render_mode": mode_desc,
                       "l2": "flushed between timed steps (256 MiB memset, untimed)"},
            "e2e": {"value": e2e_value, "unit": METRIC, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "ms_per_step": ms_e2e_step, "kind": "pipelined throughput (stream events)",
                    "wall_clock_ms_per_step_pipelined": wall_pipelined_ms, "wall_clock_value_pipelined": voice_samples / (wall_pipelined_ms * 1e-3),
                    "wall_clock_ms_single_job_latency": wall_latency_ms, "wall_clock_value_single_job": voice_samples / (wall_latency_ms * 1e-3),
                    "note": "jobs submitted back to back: a job's H2D + pre-pass (upload stream) and the previous job's D2H (copy stream) "
                            "overlap the render; every step copies its events in and its mix out.  wall_clock_*: host perf_counter over the same "
                            "K jobs without the L2 flush (pipelined) and of ONE job alone, set_events -> run -> D2H -> synchronize (latency)"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": roofline,
        }
        if args.mode == "segments":
            line["config"]["segments"] = r.segment_info()
        if parity is not None:
            line["parity"] = parity
        if strong is not None:
            line["strong"] = strong
        if sfx is not None:
            line["sfx"] = sfx
        if world == 1 and not args.no_other_modes:
            # the same job through the other render modes (device timed, inputs resident, L2 flushed), for comparison
            others = {}
            for m, mid in (("exact", L.RENDER_EXACT), ("scan", L.RENDER_SCAN), ("segments", L.RENDER_SEGMENTS)):
                if m == args.mode:
                    continue
                r2 = Render(ctx, V, total_frames, wl["color"], wl["gains"], channels=2, sample_rate=SAMPLE_RATE, buffer_frames=BUFFER_FRAMES, mode=mid)
                r2.set_events(off_host, ev_host)

                def step_other():
                    mix_t.zero_()
                    r2.run(mix)
                step_other()
                ms_o = timed(step_other, 3)
                others[m] = {"ms_per_step": sum(ms_o) / len(ms_o), "value": voice_samples / (sum(ms_o) / len(ms_o) * 1e-3), "steps": 3,
                             "render_mode": MODE_INFO[m][2]}
                r2.free()
            if args.mode == "segments":
                # opt-in variant of the benched mode: the Filter recurrence with FMAs (zcu_render_set_filter_contraction).  Not the
                # reference's bits in the Filter any more; the probe below compares it with the benched (strict) render on voice
                # windows at the segment starts of the whole job, whose own distance to the oracle is `parity`.
                r2 = Render(ctx, V, total_frames, wl["color"], wl["gains"], channels=2, sample_rate=SAMPLE_RATE, buffer_frames=BUFFER_FRAMES,
                            mode=L.RENDER_SEGMENTS)
                r2.set_events(off_host, ev_host)
                r2.set_filter_contraction(True)

                def step_contracted():
                    mix_t.zero_()
                    r2.run(mix)
                step_contracted()
                ms_o = timed(step_contracted, 3)
                probe_err, probe_n = 0.0, 0
                nwin = 2048
                for first in (0, total_frames // 3, total_frames - nwin):
                    for v0 in (0, V // 2, V - 32):
                        wa, wb = ctx.buf(nwin, 32), ctx.buf(nwin, 32)
                        r.run_voice_window(v0, first, wa)
                        r2.run_voice_window(v0, first, wb)
                        probe_err = max(probe_err, float(np.abs(wa.download().astype(np.float64) - wb.download()).max()))
                        probe_n += nwin * 32
                        wa.free()
                        wb.free()
                others["segments_contracted_filter"] = {
                    "ms_per_step": sum(ms_o) / len(ms_o), "value": voice_samples / (sum(ms_o) / len(ms_o) * 1e-3), "steps": 3,
                    "render_mode": "segments with zcu_render_set_filter_contraction(1): the Filter's multiply-add pairs as FMAs (10 instead of "
                                   "16 instructions per sample); opt-in, Filter output within the mode's 1e-5 but not bit-exact anywhere",
                    "max_abs_diff_vs_benched_render": probe_err, "voice_samples_probed": probe_n}
                r2.free()
            line["other_modes"] = others
            # The same bank through the DROP-IN module path: the reference's paint body call by call (zero, PulseOsc,
            # multiplyWithScalar, zero, Filter, zero, Envelope, multiply = 8 C-ABI calls per block over all voices,
            # examples/modules.zig:220-246), temps in HBM, on the first 64 mix buffers of the job.  This is the HBM-bound form of
            # the workload: 64 algorithmic bytes per voice-sample against the measured copy peak.
            try:
                from zang_b200.events import BankTrigger
                from zang_b200 import engine as E
                bf = 64 * BUFFER_FRAMES
                nb = E.NiceInstrumentBank(ctx, V, 0.3)
                # a steady-state block of the benched job: buffers [1024, 1088) (21.8 s into the 60 s), every voice holding the
                # note it holds there; the preceding 64 buffers are painted first so that the module states are mid-render too
                first = min(1024 * BUFFER_FRAMES, max(bf, (total_frames // bf - 1) * bf))
                wl64 = synth.window_bank(wl, first - bf, first + bf)
                trig = BankTrigger(wl64["event_off"], wl64["events"], 2 * bf, BUFFER_FRAMES)
                rows = np.zeros(wl64["events"].size, dtype=E.NiceInstrumentBank.Params)
                rows["sample_rate"], rows["freq"], rows["note_on"] = SAMPLE_RATE, wl64["events"]["freq"], wl64["events"]["note_on"]
                bufs = [ctx.buf(bf, V) for _ in range(3)]
                ctx.set_paint_mode(L.PAINT_TOLERANCE)
                sp0 = ctx.spans(V, bf, *trig.spans(0, bf))
                nb.paint(sp0, [bufs[0]], bufs[1:], rows)
                sp0.free()
                soff, sspans = trig.spans(bf, bf)
                sp = ctx.spans(V, bf, soff, sspans)
                dp = {}
                for pm, pname, ovr in ((L.PAINT_TOLERANCE, "tolerance", False), (L.PAINT_TOLERANCE, "tolerance_overwrite", True), (L.PAINT_EXACT, "exact", False)):
                    ctx.set_paint_mode(pm)

                    def step_dropin():
                        nb.paint(sp, [bufs[0]], bufs[1:], rows, overwrite=ovr)
                    l0 = ctx.launch_count
                    step_dropin()
                    calls = ctx.launch_count - l0
                    ms_d = timed(step_dropin, 3)
                    ms_avg = sum(ms_d) / len(ms_d)
                    gbs = V * bf * 64 / (ms_avg * 1e-3) / 1e9
                    dp[pname] = {"ms_per_block": ms_avg, "value": V * bf / (ms_avg * 1e-3), "algorithmic_gbs": gbs, "frac_of_hbm_peak": gbs / peak,
                                 "kernel_launches_per_block": int(calls)}
                ctx.set_paint_mode(L.PAINT_EXACT)
                line["dropin_module_path"] = {"block_frames": bf, "calls_per_block": 8, "bytes_per_voice_sample": 64, "paint_modes": dp,
                                              "block": f"mix buffers [{first // BUFFER_FRAMES}, {first // BUFFER_FRAMES + 64}) of the benched job (steady state: every voice "
                                                       "playing; round 1 measured the job's first 64 buffers, in which voices start one by one)",
                                              "span_coverage": float((sspans["end"].astype(np.int64) - sspans["start"]).sum()) / (V * bf),
                                              "note": "tolerance = Filter in time segments whose warm-up windows are computed from the spans' constants (<= 1e-5), "
                                                      "everything else bit-exact; tolerance_overwrite = the same with every `zero; paint` pair of the body as ONE "
                                                      "call in ZCU_OUT_OVERWRITE mode (5 C-ABI calls, same bits; still counted at the reference chain's 64 B); "
                                                      "exact = every module bit-exact"}
                for b in bufs + [sp]:
                    b.free()
            except Exception as e:  # the headline number must not depend on the comparison leg
                line["dropin_module_path"] = {"error": repr(e)}
            if not args.no_extra:
                line["other_configs"] = other_configs_leg(ctx, torch, timed, peak)
        if world == 1 and not args.no_cpu_baseline:
            cb = cpu_reference_run(1, 0)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        emit(line)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
