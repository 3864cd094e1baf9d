#!/usr/bin/env python3
"""Summary JSON of one `ncu --set full` capture of the batch kernel (the r0N_*_ncu_full_summary.json files).

usage: ncu_summary.py REPORT.ncu-rep STREAMS COMPRESSED_BYTES DECOMPRESSED_BYTES "what" "command" > summary.json"""
import csv, io, json, subprocess, sys

rep, streams, cbytes, dbytes, what, command = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), sys.argv[5], sys.argv[6]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
names, units, vals = rows[0], rows[1], rows[-1]
keep = ("dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum", "l1tex__", "launch__block_size",
        "launch__grid_size", "launch__registers_per_thread", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct",
        "sm__throughput.avg.pct", "sm__warps_active.avg.per_cycle_active", "smsp__average_warps_issue_stalled",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct", "smsp__thread_inst_executed_per_inst_executed.ratio")
m = {}
for n, u, v in zip(names, units, vals):
    if any(n.startswith(k) for k in keep) and "per_issue_active" in n or n in keep or any(n.startswith(k) for k in keep[2:4] + keep[7:11] + keep[12:]):
        m[n] = {"value": v, "unit": u}
def num(n):
    v, u = float(m[n]["value"]), m[n]["unit"]
    return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "Tbyte": 1e12}.get(u, 1)
traffic = num("dram__bytes_read.sum") + num("dram__bytes_write.sum")
print(json.dumps({"what": what, "command": command, "streams": streams, "decompressed_bytes": dbytes,
                  "compressed_bytes": cbytes, "algorithmic_bytes": cbytes + dbytes, "dram_traffic_bytes": traffic,
                  "traffic_over_algorithmic": traffic / (cbytes + dbytes),
                  "warp_instructions_per_output_byte": float(m["smsp__inst_executed.sum"]["value"]) / dbytes,
                  "metrics": m}, indent=1))
