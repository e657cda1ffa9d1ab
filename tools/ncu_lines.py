#!/usr/bin/env python
"""Join an ncu SASS source page (ncu -i X.ncu-rep --page source --csv --kernel-name regex:K) with nvdisasm line info of
the built library, and print per-source-line sample / instruction totals.  Developer tooling, not product code.

usage: tools/ncu_lines.py <ncu_source.csv> <mangled-kernel-substring> [top_n]
"""
import collections
import csv
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def line_table(kernel_sub):
    work = os.path.join(ROOT, "gpurun_out", "cub")
    os.makedirs(work, exist_ok=True)
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "compressed_image_b200", "libb2g.so")], cwd=work,
                   check=True, capture_output=True)
    cubin = [f for f in os.listdir(work) if f.endswith(".cubin")][0]
    dis = subprocess.run(["nvdisasm", "--print-line-info", cubin], cwd=work, capture_output=True, text=True).stdout
    table, cur, active = {}, None, False
    for ln in dis.splitlines():
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
        if m:
            active = kernel_sub in m.group(1)
            continue
        if not active:
            continue
        m = re.match(r'\s*//## File "(.*?)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            table[int(m.group(1), 16)] = (cur, m.group(2).strip())
    return table


def main():
    path, ksub = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    rows = list(csv.reader(open(path)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    ia, isamp, iinst = hdr.index("Address"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_")]
    table = line_table(ksub)
    body = []
    for r in rows[hi + 1:]:
        if r and r[0] in ("Address", "Kernel Name"):
            break  # a second launch follows
        if len(r) > iinst and r[ia]:
            body.append(r)
    base = min(int(r[ia], 16) if r[ia].startswith("0x") else int(r[ia]) for r in body)
    agg = collections.defaultdict(lambda: [0, 0, collections.Counter()])
    tot_s = tot_i = 0
    for r in body:
        a = (int(r[ia], 16) if r[ia].startswith("0x") else int(r[ia])) - base
        key = table.get(a, (None, "?"))[0]
        s, n = int(r[isamp] or 0), int(r[iinst] or 0)
        agg[key][0] += s
        agg[key][1] += n
        for i in stall_cols:
            v = int(r[i] or 0)
            if v:
                agg[key][2][hdr[i][6:]] += v
        tot_s += s
        tot_i += n
    print("total samples %d, warp instructions %d" % (tot_s, tot_i))
    src_cache = {}
    for key, (s, n, st) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        text = ""
        if key:
            f = os.path.join(ROOT, "compressed_image_b200", "csrc", key[0])
            if f not in src_cache and os.path.exists(f):
                src_cache[f] = open(f).read().splitlines()
            if f in src_cache and key[1] <= len(src_cache[f]):
                text = src_cache[f][key[1] - 1].strip()[:90]
        stalls = " ".join("%s:%d" % kv for kv in st.most_common(3))
        print("%5.1f%% smp %5.1f%% inst  %-22s %s\n        [%s]" % (100.0 * s / max(tot_s, 1), 100.0 * n / max(tot_i, 1),
              "%s:%d" % key if key else "?", text, stalls))


if __name__ == "__main__":
    main()
