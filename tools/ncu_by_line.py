#!/usr/bin/env python
"""Attribute ncu per-SASS-instruction counters to source lines (no GPU needed).

  python tools/ncu_by_line.py <report.ncu-rep> <lib.so> <kernel-substring> [top]

ncu's `--page source --print-source sass` gives per-instruction counters; nvdisasm -g gives
the file:line (innermost inline frame) of each instruction.  Joined by instruction offset.
"""
import collections
import csv
import io
import re
import subprocess
import sys
import tempfile
import os


def main():
    rep, lib, kern = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    ci = {n: hdr.index(n) for n in ("Address", "Source", "# Samples", "Instructions Executed")}
    insts = []
    for r in rows[hdr_i + 1:]:
        if len(r) < len(hdr) or not r[0].startswith("0x"):
            continue
        insts.append((int(r[ci["Address"]], 16), r[ci["Source"]].strip(), int(r[ci["# Samples"]] or 0),
                      int(r[ci["Instructions Executed"]] or 0)))
    base = insts[0][0]
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
    # every .text section whose name holds the substring (one per template instance); the one with
    # as many instructions as the report's kernel is the instance that was profiled
    cands = []
    for f in os.listdir(tmp):
        if not f.endswith(".cubin"):
            continue
        dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], capture_output=True,
                             text=True).stdout.splitlines()
        inside, cur, sec = False, ("?", 0), None
        for ln in dis:
            if ln.startswith("//---") and ".text." in ln:
                inside = kern in ln
                if inside:
                    sec = {}
                    cands.append(sec)
                continue
            if not inside:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
            if m:
                cur = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            m = re.match(r"\s*/\*([0-9a-f]+)\*/\s+(\S.*);", ln)
            if m:
                sec[int(m.group(1), 16)] = cur
    exact = [c for c in cands if len(c) == len(insts)]
    lines = exact[0] if exact else (min(cands, key=lambda c: abs(len(c) - len(insts))) if cands else {})
    if not exact:
        print("warning: no section with exactly %d instructions (%s)" % (len(insts), [len(c) for c in cands]))
    agg = collections.defaultdict(lambda: [0, 0])
    tot_s = tot_i = 0
    for addr, _src, smp, ex in insts:
        key = lines.get(addr - base, ("?", 0))
        agg[key][0] += smp
        agg[key][1] += ex
        tot_s += smp
        tot_i += ex
    dump = os.environ.get("HB_DUMP_LINE")          # "file:line": list that line's SASS with counts
    if dump:
        df, dl = dump.split(":")
        for addr, src, smp, ex in insts:
            if lines.get(addr - base, ("?", 0)) == (df, int(dl)):
                print("    %5x  ex %9d  smp %4d  %s" % (addr - base, ex, smp, src))
    print("total samples %d, warp instructions %d, sass instructions %d" % (tot_s, tot_i, len(insts)))
    byfile = collections.defaultdict(lambda: [0, 0])
    for (f, l), (s, e) in agg.items():
        byfile[f][0] += s
        byfile[f][1] += e
    for f, (s, e) in sorted(byfile.items(), key=lambda x: -x[1][0]):
        print("  %-24s samples %5.1f%%  inst %5.1f%%" % (f, 100.0 * s / tot_s, 100.0 * e / tot_i))
    print("top lines by stall samples:")
    for (f, l), (s, e) in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
        print("  %-24s:%-4d samples %5.1f%%  inst %5.1f%%" % (f, l, 100.0 * s / tot_s, 100.0 * e / tot_i))


if __name__ == "__main__":
    main()
