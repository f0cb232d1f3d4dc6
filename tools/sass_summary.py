#!/usr/bin/env python
"""SASS opcode summary of the built library: per kernel, the counts of the instructions that prove the
Blackwell-native path (tcgen05 MMA = UTCHMMA, tensor-memory loads = LDTM, TMA loads / stores = UTMALDG /
UTMASTG, tcgen05.commit / barriers = UTCBAR, cp.async = LDGSTS, programmatic dependent launch = ACQBULK /
PREEXIT) next to the total.  No GPU needed:  python tools/sass_summary.py > profiles/r02_sass_summary.md"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "hanabi_agents_b200", "_lib")
OPS = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTCBAR", "UTMACMDFLUSH", "SYNCS", "LDGSTS",
       "ACQBULK", "PREEXIT", "HMMA", "LDSM", "ATOM", "RED", "MATCH", "SHFL", "POPC", "LOP3", "IMAD"]
print("# SASS opcode summary (sm_100a), `cuobjdump -sass` of hanabi_agents_b200/_lib/*.o\n")
print("| object | kernel | total | " + " | ".join(OPS) + " |")
print("|---|---|---|" + "---|" * len(OPS))
for obj in sorted(f for f in os.listdir(LIB) if f.endswith(".o")):
    out = subprocess.run(["cuobjdump", "-sass", os.path.join(LIB, obj)], capture_output=True, text=True).stdout
    name, cnt = None, None
    kernels = []
    for ln in out.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            name, cnt = m.group(1), collections.Counter()
            kernels.append((name, cnt))
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P[0-9T]+\s+)?([A-Z0-9_]+)", ln)
        if m and cnt is not None:
            cnt["__total"] += 1
            cnt[m.group(1)] += 1
    for name, cnt in kernels:
        dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
        dem = re.sub(r"\(anonymous namespace\)::", "", dem)
        dem = re.sub(r"\(.*", "", dem)[:80]
        if cnt["__total"] < 40 and not any(cnt[o] for o in OPS[:8]):
            continue
        print("| %s | `%s` | %d | %s |" % (obj, dem, cnt["__total"], " | ".join(str(cnt[o]) if cnt[o] else "" for o in OPS)))
