"""List the loops (backward branches) of a cuobjdump -sass listing with their instruction counts and op mix."""
import re, sys, collections
lines = open(sys.argv[1]).read().splitlines()
ins = []
for l in lines:
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
addr_index = {a: i for i, (a, _) in enumerate(ins)}
for i, (a, t) in enumerate(ins):
    m = re.search(r"BRA\s+(?:\S+,\s*)?0x([0-9a-f]+)", t)
    if m:
        tgt = int(m.group(1), 16)
        if tgt < a and tgt in addr_index:
            body = ins[addr_index[tgt]:i + 1]
            ops = collections.Counter()
            for _, x in body:
                x = re.sub(r"^@!?U?P\d+\s+", "", x)
                ops[x.split()[0].split(".")[0]] += 1
            print(f"loop {tgt:#06x}..{a:#06x}: {len(body)} instrs  " + " ".join(f"{k}:{v}" for k, v in ops.most_common(14)))
