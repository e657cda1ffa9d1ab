#!/usr/bin/env python
"""SASS of a kernel with source line info (developer tooling).
usage: tools/sass_lines.py <mangled-kernel-substring> [file:lo-hi]   -> prints the instructions attributed to those lines"""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
work = "/tmp/b2g_cub"
os.makedirs(work, exist_ok=True)
for f in os.listdir(work):
    os.remove(os.path.join(work, f))
subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "compressed_image_b200", "libb2g.so")], cwd=work, check=True, capture_output=True)
cubin = [f for f in os.listdir(work) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "--print-line-info", cubin], cwd=work, capture_output=True, text=True).stdout
sel = None
if len(sys.argv) > 2:
    f, r = sys.argv[2].split(":")
    lo, hi = r.split("-")
    sel = (f, int(lo), int(hi))
active, cur, n = False, None, 0
for ln in dis.splitlines():
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
    if m:
        active = sys.argv[1] in m.group(1)
        continue
    if not active:
        continue
    m = re.match(r'\s*//## File "(.*?)", line (\d+)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m:
        n += 1
        if sel is None or (cur and cur[0] == sel[0] and sel[1] <= cur[1] <= sel[2]):
            print("%05x %-22s %s" % (int(m.group(1), 16), "%s:%d" % cur if cur else "", m.group(2).strip()))
    elif re.match(r"\s*\.L_", ln) and sel is None:
        print(ln.strip())
print("total instructions in kernel:", n, file=sys.stderr)
