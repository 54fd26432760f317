#!/usr/bin/env python
"""Join ncu's per-SASS-instruction stall samples with nvdisasm line info and aggregate per source line / function.

usage: stall_by_line.py <ncu source-page csv> <nvdisasm -g -c output> <kernel mangled-name substring> [topN]
  ncu -i X.ncu-rep --page source --csv > src.csv
  cuobjdump -xelf all librmpc.so && nvdisasm -g -c *.cubin > sass_lines.txt
"""
import csv
import re
import sys
from collections import defaultdict

src_csv, sass_txt, kern = sys.argv[1:4]
topn = int(sys.argv[4]) if len(sys.argv) > 4 else 40

# address -> (file, line)
addr2line = {}
inside = False
cur = ("?", 0)
for ln in open(sass_txt):
    if ln.startswith(".text.") and kern in ln:
        inside = True
        continue
    if inside and ln.startswith("//-----") or (inside and ln.startswith("\t.section")):
        inside = False
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*);", ln)
    if m:
        addr2line[int(m.group(1), 16)] = cur

rows = list(csv.reader(open(src_csv)))
hdr = rows[1]
ia, isrc = hdr.index("Address"), hdr.index("Source")
isamp = hdr.index("# Samples")
iinst = hdr.index("Instructions Executed")
cols = {n: hdr.index(n) for n in ("stall_long_sb", "stall_short_sb", "stall_wait", "stall_math", "stall_lg", "stall_branch_resolving",
                                  "stall_no_inst", "stall_selected", "stall_not_selected", "stall_mio", "stall_barrier")}
base = None
per_line = defaultdict(lambda: defaultdict(float))
tot = defaultdict(float)
for r in rows[2:]:
    if len(r) <= isamp or not r[ia].strip():
        continue
    try:
        a = int(r[ia], 16) if not r[ia].isdigit() else int(r[ia])
    except ValueError:
        continue
    if base is None:
        base = a
    key = addr2line.get(a - base, ("?", 0))
    s = float(r[isamp] or 0)
    per_line[key]["samples"] += s
    per_line[key]["inst"] += float(r[iinst] or 0)
    tot["samples"] += s
    tot["inst"] += float(r[iinst] or 0)
    for n, c in cols.items():
        v = float(r[c] or 0)
        per_line[key][n] += v
        tot[n] += v

print("total samples %.0f inst %.3g" % (tot["samples"], tot["inst"]), {k: int(v) for k, v in tot.items() if k.startswith("stall")})
# function ranges (device.cuh): crude map from line -> enclosing function by scanning the source
func_of = {}
try:
    import os
    path = os.path.join(os.path.dirname(__file__), "..", "..", "robotic-mpc_b200", "csrc", "rmpc_device.cuh")
    name = "?"
    for i, ln in enumerate(open(path), 1):
        m = re.match(r"^(?:template.*\n)?RD[MN]? .*?(\w+)\(", ln)
        if m and not ln.startswith(" "):
            name = m.group(1)
        func_of[i] = name
except Exception:
    pass
per_fn = defaultdict(lambda: defaultdict(float))
for (f, l), d in per_line.items():
    fn = func_of.get(l, f) if f == "rmpc_device.cuh" else f
    for k, v in d.items():
        per_fn[fn][k] += v
print("\n== by function ==")
for fn, d in sorted(per_fn.items(), key=lambda kv: -kv[1]["samples"])[:30]:
    print("%-28s samples %6.2f%%  inst %6.2f%%  long_sb %5.1f%% short_sb %5.1f%% wait %5.1f%% sel %5.1f%%" % (
        fn, 100 * d["samples"] / tot["samples"], 100 * d["inst"] / tot["inst"], 100 * d["stall_long_sb"] / max(d["samples"], 1),
        100 * d["stall_short_sb"] / max(d["samples"], 1), 100 * d["stall_wait"] / max(d["samples"], 1),
        100 * d["stall_selected"] / max(d["samples"], 1)))
print("\n== top lines ==")
for (f, l), d in sorted(per_line.items(), key=lambda kv: -kv[1]["samples"])[:topn]:
    print("%-18s:%5d %-22s samples %5.2f%% inst %5.2f%% long_sb %5.1f%%" % (f, l, func_of.get(l, "") if f == "rmpc_device.cuh" else "",
                                                                          100 * d["samples"] / tot["samples"], 100 * d["inst"] / tot["inst"],
                                                                          100 * d["stall_long_sb"] / max(d["samples"], 1)))
