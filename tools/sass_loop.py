#!/usr/bin/env python
"""Static instruction mix of a SASS address range (a hot loop) of one kernel in a .so / .cubin.
usage: sass_loop.py <file> <kernel-substring> [lo hi] [symbols_per_iteration]
Without a range: lists the backward branches (loop candidates) of the kernel."""
import collections
import re
import subprocess
import sys

ALU = {"LOP3", "SHF", "SEL", "ISETP", "VIADD", "IADD3", "LEA", "VIMNMX", "VIMNMX3", "PRMT", "BMSK", "SGXT", "MOV", "IABS", "PLOP3", "P2R", "R2P", "FSETP", "FMNMX", "FSEL", "IADD", "UIADD3", "ULOP3", "USHF"}
FMA = {"IMAD", "FFMA", "FMUL", "FADD", "HFMA2"}
FP64 = {"DFMA", "DADD", "DMUL", "DSETP"}
XU = {"FLO", "POPC", "MUFU", "I2F", "F2I", "F2F", "BREV"}
LSU = {"LDS", "STS", "LDG", "STG", "LDL", "STL", "LDGSTS", "ATOMS", "ATOMG", "RED", "LDSM"}

so, name = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)
body = next(f for f in funcs if name in f.split("\n")[0])
print("kernel:", body.split("\n")[0][:150])
ins = []
for line in body.split("\n"):
    m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
if len(sys.argv) < 5:
    for addr, t in ins:
        m = re.search(r"BRA(?:\.\S+)*\s+(?:\S+,\s*)?(0x[0-9a-f]+)", t)
        if m and int(m.group(1), 16) < addr:
            print(f"  backward branch at {addr:#x} -> {m.group(1)}  ({(addr - int(m.group(1), 16)) // 16} instructions)")
    sys.exit(0)
lo, hi = int(sys.argv[3], 16), int(sys.argv[4], 16)
per = float(sys.argv[5]) if len(sys.argv) > 5 else 1.0
ops = collections.Counter()
pipes = collections.Counter()
for addr, t in ins:
    if lo <= addr <= hi:
        toks = t.split()
        op = toks[1] if toks[0].startswith("@") else toks[0]
        base = op.split(".")[0]
        key = base + (".W" if base == "IMAD" and (".WIDE" in op or ".HI" in op) else "")
        if base == "IMAD" and (".MOV" in op or ".IADD" in op or ".SHL" in op):
            key = op.split(".U32")[0]
        ops[key] += 1
        pipe = "alu" if base in ALU else "fma" if base in FMA else "fp64" if base in FP64 else "xu" if base in XU else "lsu" if base in LSU else "other"
        pipes[pipe] += 1
tot = sum(ops.values())
print(f"range {lo:#x}..{hi:#x}: {tot} instructions = {tot / per:.1f} per symbol")
print("  pipes:", ", ".join(f"{k} {v / per:.2f}" for k, v in pipes.most_common()))
for k, v in ops.most_common(40):
    print(f"    {k:14s} {v:4d}  {v / per:6.2f}")
