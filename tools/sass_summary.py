#!/usr/bin/env python
"""Dumps the SASS of the kernels the profiles/ directory documents and a one-line census of their memory / special
instructions:  python tools/sass_summary.py [lib.so]   ->  profiles/r02_*.sass, profiles/r02_sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "densemonoslam_b200", "lib", "libdmsodom.so")
want = {"r02_k_track_768.sass": "k_trackILi768ELi1ELb0ELb0", "r02_k_track_512.sass": "k_trackILi512ELi2ELb0ELb0", "r02_k_icp_depth.sass": "k_icp_depthILb0",
        "r02_k_populate.sass": "k_populateILb1ELb0", "r02_k_maps_pyramid.sass": "k_maps_pyramidILb1ELb0"}
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
blocks, cur = {}, None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1); blocks[cur] = []
    elif cur:
        blocks[cur].append(line)
keep = re.compile(r"^(LDG|STG|LDS|STS|LDL|STL|ATOM|RED|MUFU|SHFL|BAR|MEMBAR|NANOSLEEP|REDUX|UBLKCP|UTMA|D(ADD|MUL|FMA)|ERRBAR|CGAERRBAR|CCTL)")
summary = []
for fname, pat in want.items():
    name = next((n for n in blocks if pat in n), None)
    if not name:
        continue
    dem = subprocess.run(["cu++filt", name], capture_output=True, text=True).stdout.strip() or name
    body = [l for l in blocks[name] if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l)]
    with open(os.path.join(ROOT, "profiles", fname), "w") as f:
        f.write(f"// {dem}\n" + "\n".join(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l) for l in body) + "\n")
    ops = collections.Counter()
    for l in body:
        s = re.sub(r"^\s+/\*[0-9a-f]+\*/\s+", "", l)
        s = re.sub(r"^@!?U?P\d+\s+", "", s)
        op = s.split()[0].rstrip(";")
        if keep.match(op):
            ops[op] += 1
    summary.append(f"{fname}: {dem}: {len(body)} instructions; " + ", ".join(f"{k} {v}" for k, v in sorted(ops.items())))
open(os.path.join(ROOT, "profiles", "r02_sass_summary.txt"), "w").write("\n".join(summary) + "\n")
print("\n".join(s[:400] for s in summary))
