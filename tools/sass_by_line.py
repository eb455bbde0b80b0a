#!/usr/bin/env python
"""SASS instruction histogram by source function of the step kernel (tuning aid).
usage: tools/sass_by_line.py <obj>   (needs -lineinfo)"""
import collections, os, re, subprocess, sys, tempfile
obj = os.path.abspath(sys.argv[1])
kern = sys.argv[2] if len(sys.argv) > 2 else "bb_step_kernel"
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", obj], cwd=d, check=True, capture_output=True)
cub = [os.path.join(d, f) for f in os.listdir(d) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cub], capture_output=True, text=True).stdout.splitlines()
# map file:line -> enclosing function via a crude scan of the sources
def func_table(path):
    tab, cur = [], None
    try:
        lines = open(path).read().splitlines()
    except OSError:
        return []
    for i, l in enumerate(lines, 1):
        m = re.match(r'^(?:BBD|BB_HD|BB_HD_SUB|__device__|__global__|template|static|inline).*?\b([A-Za-z_][A-Za-z0-9_]*)\s*\(', l)
        if m and not l.strip().endswith(';'):
            cur = m.group(1)
        tab.append(cur)
    return tab
tabs = {}
hist = collections.Counter()
infunc = False
cur = ("?", 0)
for l in dis:
    if l.startswith(".text."):
        infunc = kern in l
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1), int(m.group(2)))
        continue
    if infunc and re.match(r'\s+/\*[0-9a-f]{4,}\*/', l):
        f, n = cur
        if f not in tabs:
            tabs[f] = func_table(f)
        t = tabs[f]
        fn = t[n - 1] if 0 < n <= len(t) else None
        hist[(os.path.basename(f), fn)] += 1
for (f, fn), c in hist.most_common(60):
    print(f"{c:6d} {f}:{fn}")
print(sum(hist.values()), "total")
