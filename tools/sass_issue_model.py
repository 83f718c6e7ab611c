"""Static issue-time model of the hot kernel from its SASS (no GPU needed).

Measured on B200 (profiles/r2_fp64_issue_ceiling.md): an FP64 instruction issues every max(2.19, k) cycles per scheduler,
k = number of distinct register-pair source operands that do not come from the operand-reuse cache (a DFMA with three
fresh register operands takes 3 cycles: the register file delivers one 64-bit pair per cycle per scheduler); uniform-register
and immediate operands are free; every other instruction takes one issue cycle.  The sum over the right-hand side's blocks
is the time ONE scheduler needs per warp and right-hand side when it is never short of ready warps.

    python tools/sass_issue_model.py [lib.so] [kernel regex] [lo_addr hi_addr]
"""
import re
import sys

sys.path.insert(0, __file__.rsplit("/", 1)[0])
import sass_blocks as SB

FP64 = ("DFMA", "DMUL", "DADD", "DSETP")


def model(block):
    cyc, prev, n3 = 0.0, {}, 0
    for i in block:
        t = re.sub(r"^@!?U?P\d+\s+", "", i["text"])
        parts = [p.strip() for p in (t.split(None, 1)[1] if " " in t else "").split(",")]
        if i["base"] in FP64:
            srcs = parts[2:] if i["base"] == "DSETP" else parts[1:]
            k, new = 0, {}
            for slot, s in enumerate(srcs):
                m = re.match(r"^[-|~]*\|?(R\d+)(\.reuse)?\|?", s)
                if m and not s.lstrip("-|~").startswith("UR") and m.group(1) != "RZ":
                    if prev.get(slot) != m.group(1):
                        k += 1
                    if ".reuse" in s:
                        new[slot] = m.group(1)
            prev = new
            cyc += max(2.19, k)
            n3 += k >= 3
        else:
            prev = {}
            cyc += 1.0
    return cyc, n3


def main():
    a = sys.argv[1:]
    lib = a[0] if a else "ray-trace_b200/libraytrace_b200.so"
    pat = a[1] if len(a) > 1 else r"rt_integrate_kernelILi1ELb0ELb0E"
    lo, hi = (int(a[2], 16), int(a[3], 16)) if len(a) > 3 else (0, 1 << 30)
    name, insts = SB.parse(lib, pat)[0]
    print(name[:100])
    for b in SB.blocks(insts):
        if lo <= b[0]["addr"] <= hi and len(b) >= 8:
            c, n3 = model(b)
            fp = sum(1 for i in b if i["base"] in FP64)
            print(f"  {b[0]['addr']:06x}-{b[-1]['addr']:06x} n={len(b):4d} fp64={fp:4d} stall_sum={sum(i['stall'] for i in b):5d} "
                  f"issue_model={c:7.0f} fp64_with_3_fresh_regs={n3:3d} | {b[-1]['text'][:44]}")


if __name__ == "__main__":
    main()
