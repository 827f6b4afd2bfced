#!/usr/bin/env python
"""List the loops (backward branches) of one kernel in a cuobjdump -sass listing with an instruction-class histogram
per loop body — a quick static estimate of issue slots per rollout step before spending GPU time.
usage: sass_loops.py <sass file> <function-name substring> [min_body]"""
import re
import sys
from collections import Counter

path, pat = sys.argv[1], sys.argv[2]
min_body = int(sys.argv[3]) if len(sys.argv) > 3 else 8
lines = open(path).read().split("\n")
start = next(i for i, l in enumerate(lines) if "Function :" in l and pat in l)
end = next((i for i in range(start + 1, len(lines)) if "Function :" in lines[i]), len(lines))
ins = []  # (addr, text)
rx = re.compile(r"/\*([0-9a-f]{4,})\*/\s+(.*?);")
for l in lines[start:end]:
    m = rx.search(l)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
print("function", lines[start].split(":")[1].strip()[:100], "instructions", len(ins))


def cls(t):
    t = re.sub(r"^@!?U?P\d+\s+", "", t)
    op = t.split()[0]
    return op.split(".")[0] + ("." + ".".join(op.split(".")[1:2]) if op.startswith(("MUFU", "F2F", "I2F", "F2I")) else "")


addr_index = {a: i for i, (a, _) in enumerate(ins)}
for i, (a, t) in enumerate(ins):
    m = re.search(r"\bBRA\b.*?(0x[0-9a-f]+)", t)
    if m:
        tgt = int(m.group(1), 16)
        if tgt <= a and tgt in addr_index:
            j = addr_index[tgt]
            body = ins[j:i + 1]
            if len(body) < min_body:
                continue
            c = Counter(cls(x[1]) for x in body)
            print("\nloop 0x%x..0x%x  %d instructions" % (tgt, a, len(body)))
            print("  " + "  ".join("%s:%d" % kv for kv in c.most_common()))
