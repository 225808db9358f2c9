#!/usr/bin/env python
"""Static view of a kernel's SASS: every loop (backward branch) with its length and instruction mix.

  python tools/sass_loops.py <file.cubin | lib.so> <kernel-name-substring> [min_len]

Used to judge the instruction budget of the pixel loops of k_track without a GPU (the phases are
issue-bound when trackers share the GPU, so instructions per 32-pixel step is the figure of merit).
"""
import collections
import re
import subprocess
import sys


def kernels(path):
    out = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True, check=True).stdout
    cur, name = None, None
    res = {}
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            cur = res.setdefault(name, [])
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s*/\*", line)
        if m and cur is not None:
            cur.append((int(m.group(1), 16), m.group(2).strip()))
    return res


def opcode(s):
    s = re.sub(r"^@!?U?P\d+\s+", "", s)
    return s.split()[0].split(".")[0]


def main():
    path, pat = sys.argv[1], sys.argv[2]
    min_len = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    for name, ins in kernels(path).items():
        if pat not in name:
            continue
        addr_ix = {a: i for i, (a, _) in enumerate(ins)}
        print(f"== {name}: {len(ins)} instructions")
        loops = []
        for i, (a, s) in enumerate(ins):
            m = re.search(r"\bBRA\S*\s+(?:\S+,\s*)?`?\(?\.?L?_?x?_?\w*\)?", s)
            t = re.search(r"0x([0-9a-f]+)", s) if opcode(s) == "BRA" else None
            if t:
                ta = int(t.group(1), 16)
                if ta in addr_ix and ta <= a:
                    loops.append((addr_ix[ta], i))
        for b, e in loops:
            n = e - b + 1
            if n < min_len:
                continue
            ops = collections.Counter(opcode(s) for _, s in ins[b : e + 1])
            fp = ops["FFMA"] + ops["FMUL"] + ops["FADD"]
            top = ", ".join(f"{k} {v}" for k, v in ops.most_common(14))
            print(f"  loop {ins[b][0]:#06x}..{ins[e][0]:#06x}: {n} instr, fp {fp}, LDG {ops['LDG']}, LDS {ops['LDS']}, CALL {ops['CALL']}, LDL/STL {ops['LDL']}/{ops['STL']}")
            print(f"      {top}")


if __name__ == "__main__":
    main()
