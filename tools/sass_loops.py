"""Opcode mix of every loop of a kernel that contains FFMA2/FFMA work (cuobjdump -sass, no GPU needed).
usage: sass_loops.py OBJECT KERNEL_SUBSTRING [min_fma]"""
import collections
import re
import subprocess
import sys

obj, pat = sys.argv[1], sys.argv[2]
min_fma = int(sys.argv[3]) if len(sys.argv) > 3 else 8
txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)
for f in funcs[1:]:
    name = f.split("\n", 1)[0]
    if pat not in name:
        continue
    ins = []
    for l in f.split("\n"):
        m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/\s+(.*?);", l)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
    print(name[-90:], "instructions:", len(ins))
    tot = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", t).split()[0].split(".")[0] for _, t in ins)
    print("  whole kernel:", dict(tot.most_common(14)))
    for a, t in ins:
        m = re.search(r"BRA\S*\s+(?:\S+,\s*)?(0x[0-9a-f]+)", t)
        if m:
            tgt = int(m.group(1), 16)
            if tgt < a:
                body = [x for x in ins if tgt <= x[0] <= a]
                ops = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", x[1]).split()[0].split(".")[0] for x in body)
                if ops.get("FFMA2", 0) + ops.get("FFMA", 0) >= min_fma:
                    print(f"  loop {tgt:#x}-{a:#x}: {len(body)} instr:", dict(ops.most_common(12)))
    break
