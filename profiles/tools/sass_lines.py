#!/usr/bin/env python
"""SASS of a source-line range of rmpc_device.cuh, in address order, from a built library (nvdisasm line info).
    python profiles/tools/sass_lines.py lib.so FIRST LAST [--ops]"""
import collections, re, subprocess, sys, tempfile, os
lib, a, b = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=d, check=True, stdout=subprocess.DEVNULL)
cub = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, cub)], capture_output=True, text=True).stdout
cur, out = None, []
for l in txt.split("\n"):
    m = re.search(r'rmpc_device.cuh", line (\d+)', l)
    if m:
        cur = int(m.group(1)); continue
    if re.search(r'//## File', l):
        cur = None; continue
    m = re.search(r'/\*([0-9a-f]{4,6})\*/\s+(.*);', l)
    if m: out.append((cur, m.group(1), m.group(2).strip()))
idx = [i for i, (c, _, _) in enumerate(out) if c and a <= c <= b]
print("instructions attributed:", len(idx), "span", idx[0], idx[-1])
ops = collections.Counter()
for i in range(idx[0], idx[-1] + 1):
    c, ad, t = out[i]
    m = re.match(r'(@!?U?P\d+\s+)?([A-Z0-9_]+)', t)
    ops[m.group(2) if m else t] += 1
    if "--ops" not in sys.argv: print(c, ad, t)
print(ops.most_common(30))
