"""Top source lines of an ncu report by stall samples or executed instructions.

  python tools/ncu_source_top.py report.ncu-rep [N] [samples|inst]
"""
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
key = sys.argv[3] if len(sys.argv) > 3 else "samples"
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
out, cur, hdr = [], None, None
for r in csv.reader(txt.splitlines()):
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1]
    elif len(r) > 5 and r[0] == "Line No":
        hdr = r
    elif len(r) > 5 and cur and hdr:
        try:
            ln = int(r[0])
            samp = int(r[hdr.index("# Samples")] or 0)
            ins = int(r[hdr.index("Instructions Executed")] or 0)
        except ValueError:
            continue
        out.append((samp, ins, cur.split("/")[-1], ln, r[1][:100]))
out.sort(key=lambda o: o[0] if key == "samples" else o[1], reverse=True)
print("total samples", sum(o[0] for o in out), "instructions", sum(o[1] for o in out))
for o in out[:top]:
    print(o)
