#!/usr/bin/env python
"""Turn an `ncu --page raw --csv` export of the fused conv kernel's launches (one forward) into the per-launch DRAM
traffic figure that bench.py reports as roofline.traffic:
    python experiments/ncu_traffic.py gpurun_out/r01_fused_final_raw.csv profiles/r01_fused_traffic.json"""
import csv
import json
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}


def col(name):
    i = ix[name]
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1e-3, "us": 1e-6, "ns": 1e-9,
             "second": 1.0, "msecond": 1e-3, "usecond": 1e-6, "nsecond": 1e-9}[units[i]]
    return [float(r[i].replace(",", "")) * scale for r in data]


rd, wr, dur = col("dram__bytes_read.sum"), col("dram__bytes_write.sum"), col("gpu__time_duration.sum")
n = len(data)
out = {"source": sys.argv[1], "kernel": "conv1d_adain_tc_kernel", "launches": n,
       "dram_read_bytes_per_launch": sum(rd) / n, "dram_write_bytes_per_launch": sum(wr) / n,
       "traffic_bytes_per_launch": (sum(rd) + sum(wr)) / n, "ncu_time_s_per_launch": sum(dur) / n,
       "note": "ncu --set full --clock-control none, one forward at B=64 x T=320 bf16 (cold-cache, serialised launches)"}
json.dump(out, open(sys.argv[2], "w"), indent=1)
print(json.dumps(out))
