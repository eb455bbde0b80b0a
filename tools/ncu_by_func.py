#!/usr/bin/env python
"""Executed warp-instructions and stall samples of the first profiled launch, by source function.
usage: tools/ncu_by_func.py file.ncu-rep [top_n]   (kernel must be built with -lineinfo, capture with --import-source on)"""
import collections, csv, io, os, re, subprocess, sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout


def func_table(path):
    local = path if os.path.exists(path) else os.path.join(ROOT, path.split("/root/repo/")[-1])
    tab, cur = [], None
    try:
        lines = open(local).read().splitlines()
    except OSError:
        return []
    for l in lines:
        m = re.match(r'^(?:BBD|BB_HD|BB_HD_SUB|__device__|__global__|template|static|inline)[^;]*?\b([A-Za-z_][A-Za-z0-9_]*)\s*\(', l)
        if m and not l.strip().endswith(';'):
            cur = m.group(1)
        tab.append(cur)
    return tab


path, hdr, tabs = None, None, {}
inst, samp, lines_ = collections.Counter(), collections.Counter(), collections.Counter()
launch_seen = 0
for row in csv.reader(io.StringIO(raw)):
    if not row:
        continue
    if row[0] == "File Path":
        path = row[1]
        continue
    if row[0] == "Function Name":
        continue
    if row[0] == "Kernel Name":
        launch_seen += 1
        if launch_seen > 1:
            break
        continue
    if row[0] == "Line No":
        hdr = row
        i_inst, i_samp = hdr.index("Instructions Executed"), hdr.index("# Samples")
        continue
    if hdr is None or not row[0].isdigit():
        continue
    n = int(row[0])
    if path not in tabs:
        tabs[path] = func_table(path)
    t = tabs[path]
    fn = t[n - 1] if 0 < n <= len(t) else None
    key = (os.path.basename(path), fn)
    try:
        inst[key] += int(row[i_inst]); samp[key] += int(row[i_samp])
        lines_[(os.path.basename(path), n, row[1].strip()[:70])] += int(row[i_samp])
    except ValueError:
        pass
ti, ts = sum(inst.values()) or 1, sum(samp.values()) or 1
print(f"{'inst':>10} {'%':>6} {'samples':>8} {'%':>6}  function")
for k, v in samp.most_common(top):
    print(f"{inst[k]:10d} {100 * inst[k] / ti:6.1f} {v:8d} {100 * v / ts:6.1f}  {k[0]}:{k[1]}")
print(f"{ti:10d}        {ts:8d}         total")
print("\nhottest source lines by stall samples:")
for k, v in lines_.most_common(25):
    print(f"{v:6d} {100 * v / ts:5.1f}%  {k[0]}:{k[1]}  {k[2]}")
