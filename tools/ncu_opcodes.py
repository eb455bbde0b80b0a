#!/usr/bin/env python
"""Executed-instruction picture of the first profiled launch of an .ncu-rep (captured with --import-source on):
per-warp executed instructions, opcode histogram, stall reasons, and the executed SASS listing with samples.
usage: tools/ncu_opcodes.py file.ncu-rep [listing.txt]"""
import collections, csv, io, subprocess, sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]
iS, iN, iE = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
opc, stalls = collections.Counter(), collections.Counter()
listing, warps = [], None
for r in rows[2:]:
    if len(r) < len(hdr) or not r[0].startswith("0x"):
        continue
    e, sm = int(r[iE]), int(r[iN])
    if warps is None:
        warps = e
    if e == 0:
        continue
    src = r[iS].strip()
    tok = src.split()
    op = tok[1] if tok[0].startswith("@") else tok[0]
    opc[op.split(".")[0]] += e
    st = {h[6:]: int(r[i]) for i, h in stall_cols if r[i] not in ("", "0")}
    for k, v in st.items():
        stalls[k] += v
    listing.append((e, sm, src, st))
tot = sum(opc.values())
print(f"warps {warps}; executed {tot} = {tot / warps:.0f} per warp; samples {sum(l[1] for l in listing)}")
print("opcodes per warp:", {k: round(v / warps, 1) for k, v in opc.most_common(45)})
print("stall samples:", dict(stalls.most_common()))
if len(sys.argv) > 2:
    with open(sys.argv[2], "w") as f:
        n = 0
        for e, sm, src, st in listing:
            n += e / warps
            f.write(f"{n:7.0f} {e:6d} {sm:3d} {src:64s} {' '.join(f'{k}={v}' for k, v in st.items())}\n")
