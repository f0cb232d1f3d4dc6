#!/usr/bin/env python
"""Static SASS instruction mix of one kernel by issue pipe (no GPU needed).

  python tools/sass_mix.py <obj-or-so> <kernel-substring> [<second-substring>]

ALU pipe: LOP3 SHF SEL ISETP IADD3 VIADD LEA PRMT VIMNMX* PLOP3 (one warp instruction per two cycles
per SM sub-partition); FMA pipe: IMAD* FFMA FMUL FADD (same rate, issues beside the ALU pipe).
"""
import collections, re, subprocess, sys

ALU = ("LOP3", "SHF", "SEL", "ISETP", "IADD3", "VIADD", "LEA", "PRMT", "VIMNMX", "VIMNMX3", "VIADDMNMX",
       "PLOP3", "FSETP", "FSEL", "FMNMX", "IABS", "BMSK", "SGXT", "FLO", "BREV", "POPC", "MOV", "IADD")
FMA = ("IMAD", "FFMA", "FMUL", "FADD", "HFMA2", "HADD2", "HMUL2")

def main():
    obj, k1 = sys.argv[1], sys.argv[2]
    k2 = sys.argv[3] if len(sys.argv) > 3 else ""
    out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    inside = False
    ins = []
    for ln in out.splitlines():
        if "Function :" in ln:
            inside = k1 in ln and k2 in ln
            continue
        if not inside:
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(?:@!?U?P[0-9T]+\s+)?([A-Z0-9_]+)((?:\.[A-Z0-9_]+)*)(.*)", ln)
        if m:
            ins.append((int(m.group(1), 16), m.group(2), m.group(4)))
    # the widest backward branch = the step loop of the fused kernels
    lo, hi = 0, 1 << 60
    best = 0
    for addr, op, rest in ins:
        if op == "BRA":
            t = re.search(r"0x([0-9a-f]+)", rest)
            if t and int(t.group(1), 16) < addr and addr - int(t.group(1), 16) > best:
                best = addr - int(t.group(1), 16)
                lo, hi = int(t.group(1), 16), addr
    if "--all" in sys.argv or best == 0:
        lo, hi = 0, 1 << 60
    print("range 0x%x..0x%x of %d instructions" % (lo, min(hi, ins[-1][0]), len(ins)))
    ops = collections.Counter(op for addr, op, _ in ins if lo <= addr <= hi)
    alu = sum(v for k, v in ops.items() if k in ALU)
    fma = sum(v for k, v in ops.items() if k in FMA)
    tot = sum(ops.values())
    print("total %d  alu %d  fma %d  other %d" % (tot, alu, fma, tot - alu - fma))
    print(" ".join("%s:%d" % kv for kv in ops.most_common(24)))

main()
