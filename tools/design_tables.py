"""Prints the §4.5 tables of DESIGN.md from the measurement files under profiles/ (tag given on the command line).
  python tools/design_tables.py r01d [r01b]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
old = sys.argv[2] if len(sys.argv) > 2 else "r01b"
P = lambda n: json.load(open(os.path.join(ROOT, "profiles", n)))
a = P(f"{tag}_module_bw_4096.json")["results"]
b = {r["module"]: r for r in P(f"{tag}_module_bw_65536.json")["results"]}
o = {r["module"]: r for r in P(f"{old}_module_bw_65536.json")["results"]}
TP = {"SineOsc const": "exact skip-ahead", "PulseOsc const": "exact skip-ahead", "TriSawOsc const": "exact skip-ahead",
      "PulseOsc freq buf": "exact chained integer scan", "Noise white": "exact GF(2) jump", "Noise pink": "exact GF(2) jump + filter replay", "Envelope": "exact skip-ahead",
      "Gate": "stateless", "Filter const": "tolerance: verified warm-up / chain", "Filter cutoff+res buf": "tolerance: verified warm-up / chain",
      "Portamento": "exact skip-ahead", "Decimator": "exact (prefix table of one counter replay per voice)", "Distortion overdrive": "stateless",
      "Distortion clip": "stateless", "Cycle const": "exact skip-ahead", "Curve smoothstep": "no (per-call node bookkeeping)",
      "Sampler resample x0.73": "exact skip-ahead", "SimpleDelay 7500": "no (the line is carried state)",
      "FilteredEchoes 15000": "no (feedback through the line)", "FmOscillator freq buf": "no (nonlinear feedback recurrence)"}
print(f"| module (variant) | B / voice-sample | 4096 voices x 64 buffers: GB/s (of peak) | 65536 voices x 8 buffers: GB/s (of peak) | {old} | time-parallel |")
print("|---|---|---|---|---|---|")
for r in a:
    m = r["module"]
    if m.startswith("basics"):
        continue
    x = b[m]
    print(f"| {m} | {r['bytes_per_voice_sample']} | {r['gbs']:.0f} ({r['frac_of_peak']:.2f}) | {x['gbs']:.0f} ({x['frac_of_peak']:.2f}) | {(f"{o[m]['frac_of_peak']:.2f}" if m in o else "-")} | {TP.get(m, 'no (one segment)')} |")
print()
print("basics @65536:", ", ".join(f"{m.split('.')[1]} {r['gbs']:.0f}" for m, r in b.items() if m.startswith("basics")))
print()
print(f"| config (drop-in module path, `ZCU_PAINT_TOLERANCE`) | voices x frames | launches / block | ms | voice-samples/s | algorithmic GB/s (of peak) | {old} | host span tables |")
print("|---|---|---|---|---|---|---|---|")
for r, q in zip(P(f"{tag}_bank_bench.json")["results"], P(f"{old}_bank_bench.json")["results"]):
    print(f"| {r['config']} | {r['voices']} x {r['frames']} | {r['launches_per_block']} | {r['ms']:.2f} | {r['voice_samples_per_s']:.2e} | {r['gbs']:.0f} ({r['frac_of_peak']:.2f}) | {q['frac_of_peak']:.2f} | {r['host_span_tables_ms']:.1f} ms |")
