#!/usr/bin/env python
"""Per-kernel resource and SASS summary of libbball.so (cuobjdump): registers, shared memory, stack / spill bytes, SASS
instruction count, and the tensor-core / bulk-copy opcode mix.   usage: tools/sass_summary.py [lib.so] > profiles/sass_summary.txt"""
import collections, os, re, subprocess, sys

lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                           "madrona_basketball_b200", "libbball.so")
res = subprocess.run(["cuobjdump", "-res-usage", lib], capture_output=True, text=True).stdout
usage = {}
for m in re.finditer(r"Function (\S+):\n\s+(.*)", res):
    usage[m.group(1)] = dict(kv.split(":") for kv in m.group(2).split() if ":" in kv)
elf = subprocess.run(["cuobjdump", "-lelf", lib], capture_output=True, text=True).stdout.split()
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
count, ops = collections.Counter(), collections.defaultdict(collections.Counter)
name = None
WATCH = ("HMMA", "IMMA", "DMMA", "UTCHMMA", "UTCQMMA", "UTCMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UBLKPF", "LDGSTS",
         "SYNCS", "MUFU", "DFMA", "LDS", "STS", "LDG", "STG", "SHFL", "BAR", "ATOM", "RED", "CALL")
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = m.group(1)
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and name:
        count[name] += 1
        op = m.group(1)
        for w in WATCH:
            if op.startswith(w):
                ops[name][w] += 1
                break
demangle = subprocess.run(["c++filt"] + list(count), capture_output=True, text=True).stdout.splitlines()
print(f"# {os.path.basename(lib)}: ELF images {', '.join(sorted(set(e for e in elf if 'sm_' in e)))}")
print("# static SASS (every path, hot or cold); REG = registers / thread, SHARED = static shared bytes, STACK / spills from ptxas")
print(f"{'instr':>7} {'REG':>4} {'SHARED':>7} {'STACK':>6}  kernel  |  opcode mix")
for (mangled, n), pretty in sorted(zip(count.items(), demangle), key=lambda x: -x[0][1]):
    u = usage.get(mangled, {})
    short = re.sub(r"\(.*", "", pretty).replace("void ", "").replace("bb::", "")
    mix = " ".join(f"{k}={v}" for k, v in sorted(ops[mangled].items(), key=lambda x: -x[1]))
    print(f"{n:7d} {u.get('REG', '?'):>4} {u.get('SHARED', '?'):>7} {u.get('STACK', '?'):>6}  {short}  |  {mix}")
