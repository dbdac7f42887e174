"""Static instruction count of the per-sample code of every module inside a fused graph kernel (no GPU needed):
builds the one-module graph `zero t0; m.paint -> t0; out += t0`, compiles it with NVRTC, disassembles the cubin and
prints the size of the unrolled (x4) all-live loop divided by 4."""
import os, re, subprocess, sys, tempfile, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
d = tempfile.mkdtemp()
os.environ["ZCU_GRAPH_DUMP"] = d
from zang_b200 import engine as E, _lib as L  # noqa: E402

CASES = [("SineOsc const", L.MOD_SINE_OSC, ()), ("SineOsc freq+phase buf", L.MOD_SINE_OSC, (E.IN(0), E.IN(1))),
         ("PulseOsc const", L.MOD_PULSE_OSC, ()), ("PulseOsc freq buf", L.MOD_PULSE_OSC, (E.IN(0),)),
         ("TriSawOsc const", L.MOD_TRISAW_OSC, ()), ("TriSawOsc freq buf", L.MOD_TRISAW_OSC, (E.IN(0),)),
         ("Noise", L.MOD_NOISE, ()), ("Envelope", L.MOD_ENVELOPE, ()), ("Gate", L.MOD_GATE, ()),
         ("Filter const", L.MOD_FILTER, (E.IN(0),)), ("Filter cutoff buf", L.MOD_FILTER, (E.IN(0), E.IN(1))),
         ("Portamento", L.MOD_PORTAMENTO, ()), ("Curve", L.MOD_CURVE, ()), ("Sampler", L.MOD_SAMPLER, ()),
         ("Decimator", L.MOD_DECIMATOR, (E.IN(0),)), ("Distortion", L.MOD_DISTORTION, (E.IN(0),)), ("Cycle const", L.MOD_CYCLE, ())]


def loops(sass):
    ins = []
    for l in sass.splitlines():
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
    idx = {a: i for i, (a, _) in enumerate(ins)}
    out = []
    for i, (a, t) in enumerate(ins):
        m = re.search(r"BRA\s+(?:\S+,\s*)?0x([0-9a-f]+)", t)
        if m and int(m.group(1), 16) < a and int(m.group(1), 16) in idx:
            body = ins[idx[int(m.group(1), 16)]:i + 1]
            ops = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", x).split()[0].split(".")[0] for _, x in body)
            out.append((len(body), ops))
    return out


for name, kind, ins in CASES:
    b = E.GraphBuilder()
    b.zero(E.T(0))
    b.call(kind, E.T(0), ins)
    b.addInto(E.OUT(0), E.T(0))
    cfg = E.graph_config(4096, b, 1, num_inputs=len(ins))
    rc, log, n = E.graph_compile_check(cfg, b)
    if rc:
        print(name, "FAILED", log[:300])
        continue
    sass = subprocess.run(["cuobjdump", "-sass", os.path.join(d, "zcu_graph_kernel.cubin")], capture_output=True, text=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", os.path.join(d, "zcu_graph_kernel.cubin")], capture_output=True, text=True).stdout
    reg = re.search(r"REG:(\d+) STACK:(\d+)", res)
    ls = [x for x in loops(sass) if x[1].get("STG", 0) >= 4]  # the unrolled loop stores four outputs
    big = min(ls, key=lambda x: x[0]) if ls else (0, collections.Counter())
    print(f"{name:24s} regs {reg.group(1):>3s} stack {reg.group(2):>4s}  unrolled loop {big[0]:5d} instrs = {big[0] / 4:6.1f} per sample   "
          + " ".join(f"{k}:{v}" for k, v in big[1].most_common(8)))
