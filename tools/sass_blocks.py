"""Static issue-time model of a kernel from its SASS control words (no GPU needed).

For every basic block of one function: instruction count, FP64-pipe instructions (DFMA/DMUL/DADD/DSETP + MUFU.*64H
feeders are listed separately) and the sum of the control words' stall counts = the cycles ONE warp needs to issue
the block when nothing else delays it (fixed-latency dependencies are encoded in the stall counts; variable-latency
ones — loads, MUFU — wait on scoreboard barriers and are not included).  With W warps per scheduler the block cannot
run faster than max(stall_sum / W, 2 * fp64 (one FP64 warp instruction per 2 cycles per scheduler), instructions).

    python tools/sass_blocks.py [lib.so] [regex on the function name] [--min N]
"""
import re
import subprocess
import sys

FP64 = ("DFMA", "DMUL", "DADD", "DSETP")


def parse(lib, pattern):
    txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    out = []
    for f in re.split(r"\n\s*Function : ", txt)[1:]:
        name = f.split("\n", 1)[0].strip()
        if not re.search(pattern, name):
            continue
        insts = []
        lines = f.split("\n")
        for k, ln in enumerate(lines):
            m = re.match(r"^\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s*/\* 0x([0-9a-f]{16}) \*/", ln)
            if not m:
                continue
            m2 = re.match(r"^\s+/\* 0x([0-9a-f]{16}) \*/", lines[k + 1]) if k + 1 < len(lines) else None
            hi = int(m2.group(1), 16) if m2 else 0
            ctrl = hi >> 41
            text = m.group(2).strip()
            t = re.sub(r"^@!?U?P\d+\s+", "", text)
            op = t.split()[0]
            insts.append(dict(addr=int(m.group(1), 16), text=text, op=op, base=op.split(".")[0], stall=ctrl & 15,
                              yield_=(ctrl >> 4) & 1, wr=(ctrl >> 5) & 7, rd=(ctrl >> 8) & 7, wait=(ctrl >> 11) & 63))
        out.append((name, insts))
    return out


def blocks(insts):
    targets = set()
    for i in insts:
        if i["base"] in ("BRA", "BSSY", "CALL", "BRX", "JMP", "BSYNC", "WARPSYNC", "RET", "EXIT", "BREAK"):
            for m in re.finditer(r"0x([0-9a-f]+)", i["text"]):
                targets.add(int(m.group(1), 16))
    bl, cur = [], []
    for i in insts:
        if i["addr"] in targets and cur:
            bl.append(cur)
            cur = []
        cur.append(i)
        if i["base"] in ("BRA", "CALL", "BRX", "JMP", "RET", "EXIT", "BSYNC"):
            bl.append(cur)
            cur = []
    if cur:
        bl.append(cur)
    return bl


def summarize(b):
    n = len(b)
    fp = sum(1 for i in b if i["base"] in FP64)
    mufu = sum(1 for i in b if i["base"] == "MUFU")
    stall = sum(i["stall"] for i in b)
    lds = sum(1 for i in b if i["base"] in ("LDS", "STS", "LDG", "STG", "LDL", "STL", "LDC", "LDCU"))
    waits = sum(1 for i in b if i["wait"])
    return n, fp, mufu, stall, lds, waits


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    lib = args[0] if args else "ray-trace_b200/libraytrace_b200.so"
    pat = args[1] if len(args) > 1 else r"rt_integrate_kernelILi1ELb0E"
    mn = 40
    if "--min" in sys.argv:
        mn = int(sys.argv[sys.argv.index("--min") + 1])
    for name, insts in parse(lib, pat):
        tot = summarize(insts)
        print(f"== {name[:90]}")
        print(f"   total: {tot[0]} instructions ({tot[0] * 16 / 1024:.1f} KB), fp64 {tot[1]}, mufu {tot[2]}, stall sum {tot[3]}")
        for b in blocks(insts):
            n, fp, mufu, stall, lds, waits = summarize(b)
            if n >= mn:
                last = b[-1]["text"][:40]
                print(f"   [{b[0]['addr']:06x}-{b[-1]['addr']:06x}] n={n:5d} fp64={fp:5d} mufu={mufu:3d} mem={lds:3d} "
                      f"waits={waits:3d} stall={stall:6d}  stall/inst={stall / n:4.2f}  pipe-bound={2 * fp:5d}  | {last}")


if __name__ == "__main__":
    main()
