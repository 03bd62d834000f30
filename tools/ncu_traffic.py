#!/usr/bin/env python
"""DRAM traffic per launch of the kernels in an .ncu-rep (`ncu --set full` capture), merged into profiles/traffic.json,
which bench.py reads for the `roofline.traffic` field of its JSON line.
usage: python tools/ncu_traffic.py <workload> <rep> [<rep> ...]     (kernel key = function name without template args)"""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def main():
    workload, reps = sys.argv[1], sys.argv[2:]
    path = os.path.join(ROOT, "profiles", "traffic.json")
    table = json.load(open(path)) if os.path.exists(path) else {}
    entry = table.setdefault(workload, {})
    for rep in reps:
        out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--metrics", "dram__bytes_read.sum,dram__bytes_write.sum"],
                             capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(out)))
        hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
        h, units = rows[hi], rows[hi + 1]
        kn, rd, wr = h.index("Kernel Name"), h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum")
        for r in rows[hi + 2:]:
            if len(r) <= max(rd, wr):
                continue
            name = re.sub(r"^void\s+", "", r[kn])
            name = re.sub(r"<.*", "", name).split("(")[0].split("::")[-1]
            b = float(r[rd].replace(",", "")) * UNIT[units[rd]] + float(r[wr].replace(",", "")) * UNIT[units[wr]]
            stage = next((st for st in ("histogram", "encode", "decode", "compact")
                          if name.endswith(st + "_kernel") and "prelude" not in name and "tables" not in name), None)
            if stage is None:
                continue
            entry[stage] = {"kernel": name, "dram_bytes_per_launch": b, "read": float(r[rd].replace(",", "")) * UNIT[units[rd]],
                           "write": float(r[wr].replace(",", "")) * UNIT[units[wr]], "source": os.path.basename(rep)}
    json.dump(table, open(path, "w"), indent=1, sort_keys=True)
    print(json.dumps(entry, indent=1))


if __name__ == "__main__":
    main()
