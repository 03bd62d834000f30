#!/usr/bin/env python
"""Summarise an .ncu-rep here (no GPU needed): per kernel the headline metrics, the dynamic
instruction mix by opcode, shared-memory wavefronts and the stall-reason totals.
usage: python tools/ncu_summary.py gpurun_out/x/prof.ncu-rep [n_symbols] > profiles/xyz.txt"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
nsym = float(sys.argv[2]) if len(sys.argv) > 2 else None

WANT = ["Duration", "DRAM Throughput", "Memory Throughput", "Compute (SM) Throughput", "Registers Per Thread",
        "Achieved Occupancy", "Theoretical Occupancy", "Executed Ipc Active", "Issue Slots Busy",
        "Eligible Warps Per Scheduler", "Active Warps Per Scheduler", "Warp Cycles Per Issued Instruction",
        "Executed Instructions", "L1/TEX Hit Rate", "Dynamic Shared Memory Per Block", "Grid Size", "Block Size",
        "Waves Per SM", "Mem Busy", "Max Bandwidth", "L2 Cache Throughput", "Avg. Active Threads Per Warp",
        "SM Frequency", "DRAM Frequency", "Elapsed Cycles"]


def run(args):
    return subprocess.run(["ncu", "-i", rep] + args, capture_output=True, text=True).stdout


det = list(csv.reader(io.StringIO(run(["--page", "details", "--csv"]))))
h = det[0]
ix = {k: h.index(k) for k in ("ID", "Kernel Name", "Metric Name", "Metric Value", "Metric Unit")}
kernels = collections.OrderedDict()
for r in det[1:]:
    kernels.setdefault((r[ix["ID"]], r[ix["Kernel Name"]]), []).append(r)
for (kid, name), rows in kernels.items():
    print(f"=== [{kid}] {name}")
    for r in rows:
        if r[ix["Metric Name"]] in WANT:
            print(f"    {r[ix['Metric Name']]:42s} {r[ix['Metric Value']]:>16s} {r[ix['Metric Unit']]}")

raw = list(csv.reader(io.StringIO(run(["--page", "raw", "--csv"]))))
if raw:
    h = raw[0]
    pick = [i for i, c in enumerate(h) if any(k in c for k in (
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum", "smsp__inst_executed.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__cycles_active.avg", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "lts__t_bytes.sum"))
            and "per_second" not in c and "pct" not in c.replace("sm__warps_active.avg.pct", "")]
    name_i = h.index("Kernel Name")
    for r in raw[2:]:
        print(f"--- raw: {r[name_i][:60]}")
        for i in pick:
            print(f"    {h[i]:70s} {r[i]:>18s} {raw[1][i]}")

# source page per kernel
for (kid, name) in kernels:
    short = name.split("(")[0].split("::")[-1].split("<")[0].replace("void ", "").strip()
    src = list(csv.reader(io.StringIO(run(["--page", "source", "--csv", "--kernel-name", "regex:" + short]))))
    hdr_i = next((i for i, r in enumerate(src) if r and r[0] == "Address"), None)
    if hdr_i is None:
        continue
    h = src[hdr_i]
    col = {c: i for i, c in enumerate(h)}
    ops = collections.Counter()
    wave = collections.Counter()
    stalls = collections.Counter()
    tot = 0
    seen = set()
    for r in src[hdr_i + 1:]:
        if len(r) < len(h) or not r[col["Instructions Executed"]].isdigit():
            continue
        if r[col["Address"]] in seen:  # the page repeats the SASS rows once per view
            continue
        seen.add(r[col["Address"]])
        sass = r[col["Source"]].strip()
        toks = sass.split()
        op = toks[1] if toks and toks[0].startswith("@") and len(toks) > 1 else (toks[0] if toks else "?")
        op = ".".join(op.split(".")[:2]) if op.startswith(("LD", "ST", "ATOM", "RED")) else op.split(".")[0]
        n = int(r[col["Instructions Executed"]] or 0)
        ops[op] += n
        tot += n
        w = int(r[col["L1 Wavefronts Shared"]] or 0) if "L1 Wavefronts Shared" in col else 0  # absent: no smem access
        if w:
            wave[op] += w
            wave[op + " (ideal)"] += int(r[col["L1 Wavefronts Shared Ideal"]] or 0)
        for c, i in col.items():
            if c.startswith("stall_") and "Not Issued" not in c:
                try:
                    stalls[c] += int(r[i] or 0)
                except ValueError:
                    pass
    print(f"=== instruction mix [{kid}] {short}: {tot} warp instructions" + (f" = {tot * 32 / nsym:.1f} per 32 symbols" if nsym else ""))
    for op, n in ops.most_common(28):
        print(f"    {op:14s} {n:>12d} {100.0 * n / max(tot, 1):5.1f}%" + (f"  {n * 32 / nsym:6.2f}/32sym" if nsym else ""))
    if wave:
        print("    shared-memory wavefronts:", dict(wave))
        if nsym:
            tw = sum(v for k, v in wave.items() if "ideal" not in k)
            print(f"    = {tw * 32 / nsym:.2f} wavefronts per 32 symbols")
    st = sum(stalls.values())
    print("    stall samples:", ", ".join(f"{k[6:]} {100.0 * v / max(st, 1):.1f}%" for k, v in stalls.most_common(9)))
