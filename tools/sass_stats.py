"""Static SASS instruction mix per kernel of the built library (cuobjdump -sass)."""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "ray-trace_b200/libraytrace_b200.so"
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
for f in re.split(r"\n\s*Function : ", txt)[1:]:
    name = f.split("\n", 1)[0]
    ops = collections.Counter()
    for m in re.finditer(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", f, re.M):
        ops[m.group(1).split(".")[0]] += 1
    tot = sum(ops.values())
    fp64 = sum(v for k, v in ops.items() if k in ("DFMA", "DMUL", "DADD", "DSETP", "MUFU"))
    print(f"{name[:70]:70s} total={tot:6d} ({tot * 16 // 1024} KB) fp64={fp64}", dict(ops.most_common(8)))
