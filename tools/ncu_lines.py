#!/usr/bin/env python
"""Join an ncu report's per-SASS-instruction counters with nvdisasm -g line info and aggregate per source line.

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep minigrid_dynamicprogramming_b200/libmgb200.so step_many [top] [launch_skip] [samples]

The third argument is a regex on the MANGLED kernel name and must select ONE instantiation (e.g. step_many_kernelILb0ELi0);
the second may also be a .cubin.  launch_skip picks the (launch_skip+1)-th matching launch of the report.
"""
import collections
import csv
import glob
import os
import re
import subprocess
import sys
import tempfile

rep, so, pat = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
skip = sys.argv[5] if len(sys.argv) > 5 else "0"
by_samples = len(sys.argv) > 6 and sys.argv[6] == "samples"      # order by stall samples instead of instruction count
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", skip, "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if "Instructions Executed" in r)
hdr = rows[hi]
iA, iS, iE, iSm = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
insts = []
for r in rows[hi + 1:]:
    if len(r) < len(hdr):
        break                       # next kernel
    try:
        insts.append((r[iS], int(r[iE]), int(r[iSm])))
    except ValueError:
        pass
tmp = tempfile.mkdtemp()
if so.endswith(".cubin"):
    cubins = [os.path.abspath(so)]
else:
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, capture_output=True)
    cubins = sorted(glob.glob(os.path.join(tmp, "*.cubin")))
lines = []
for cubin in cubins:
    txt = subprocess.run(["nvdisasm", "-g", cubin], capture_output=True, text=True).stdout
    infun, cur = False, None
    for ln in txt.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", ln)
        if m:
            infun = re.search(pat, m.group(1)) is not None
            continue
        if not infun:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*);", ln)
        if m:
            lines.append((cur, m.group(2).strip()))
    if lines:
        break
if len(lines) != len(insts):
    print(f"warning: {len(lines)} SASS lines in the .so vs {len(insts)} in the report (different build?)")
agg = collections.defaultdict(lambda: [0, 0])
tot = 0
for (loc, _), (_, e, s) in zip(lines, insts):
    agg[loc][0] += e
    agg[loc][1] += s
    tot += e
srcs = {}
print(f"total warp-instructions executed: {tot}")
for loc, (e, s) in sorted(agg.items(), key=lambda kv: -kv[1][1 if by_samples else 0])[:top]:
    text = ""
    if loc:
        f = glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "minigrid_dynamicprogramming_b200", "csrc", loc[0]))
        if f:
            srcs.setdefault(f[0], open(f[0]).read().splitlines())
            text = srcs[f[0]][loc[1] - 1].strip()[:90]
    print(f"{e:11d} {100 * e / tot:5.1f}%  samples {s:5d}  {loc}  {text}")
