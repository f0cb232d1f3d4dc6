#!/usr/bin/env python
"""Static SASS instruction count per source line of one kernel (no GPU needed).

  python tools/sass_by_line.py <obj-or-so> <kernel-substring> [<second-substring>]

The fused self-play kernel is close to straight-line code that a warp walks once per step, so the
static count per line is a good proxy of the executed count (profiles/r01_selfplay_history.md).
"""
import collections, os, re, subprocess, sys, tempfile

def main():
    obj, k1 = sys.argv[1], sys.argv[2]
    k2 = sys.argv[3] if len(sys.argv) > 3 else ""
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
    agg = collections.Counter(); n = 0
    for f in os.listdir(tmp):
        if not f.endswith(".cubin"): continue
        dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
        inside, cur = False, ("?", 0)
        for ln in dis.splitlines():
            if ln.startswith("//---") and ".text." in ln:
                inside = k1 in ln and k2 in ln; continue
            if not inside: continue
            m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
            if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
            if re.match(r"\s*/\*([0-9a-f]+)\*/\s+(\S.*);", ln): agg[cur] += 1; n += 1
    print("total", n)
    files = collections.Counter()
    for (f, l), v in agg.items(): files[f] += v
    print(dict(files))
    for (f, l), v in sorted(agg.items()): print(f"{f}:{l}\t{v}")

main()
