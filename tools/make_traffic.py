#!/usr/bin/env python
"""Turns one `ncu --set full` capture of a bench launch into an entry of profiles/traffic.json.

usage: ncu -i X.ncu-rep --page raw --csv | python tools/make_traffic.py <problem> <batch> <horizon> <warm|cold> "<source text>" [out.json]

bench.py reports `roofline.traffic` only when the entry's lib_sha256 equals the sha256 of the libpantr_b200.so it is
running (the capture says nothing about another build of the kernel) and problem / batch / horizon / start match the
dominant launch of the run; otherwise it prints `traffic: null`."""
import csv
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
problem, batch, horizon, start, source = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], sys.argv[5]
out = sys.argv[6] if len(sys.argv) > 6 else os.path.join(ROOT, "profiles", "traffic.json")
rows = list(csv.reader(sys.stdin))
hdr, vals = rows[0], rows[2]
col = {h: i for i, h in enumerate(hdr)}
units = rows[1]


def gb(name):
    v, u = float(vals[col[name]]), units[col[name]]
    return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}[u]


rd, wr = gb("dram__bytes_read.sum"), gb("dram__bytes_write.sum")
lib = os.path.join(ROOT, "pantr_cdc2023_experiments_b200", "libpantr_b200.so")


sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from source_hash import source_sha256  # noqa: E402


entry = {"lib_sha256": hashlib.sha256(open(lib, "rb").read()).hexdigest(), "src_sha256": source_sha256(), "problem": problem, "batch": batch, "horizon": horizon,
         "start": start, "kernel": vals[col["Kernel Name"]], "dram_bytes_per_launch": rd + wr, "dram_bytes_read": rd,
         "dram_bytes_write": wr, "kernel_ms_under_ncu": float(vals[col["gpu__time_duration.sum"]]) * {"ms": 1.0, "us": 1e-3, "s": 1e3, "ns": 1e-6}[units[col["gpu__time_duration.sum"]]],
         "source": source}
data = {"captures": []}
if os.path.exists(out):
    try:
        old = json.load(open(out))
        if "captures" in old:
            data = old
    except ValueError:
        pass
key = lambda e: (e["problem"], e["batch"], e["horizon"], e["start"])
data["captures"] = [e for e in data["captures"] if key(e) != key(entry)] + [entry]
with open(out, "w") as fh:
    json.dump(data, fh, indent=1)
    fh.write("\n")
print(json.dumps(entry))
