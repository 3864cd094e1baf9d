#!/usr/bin/env python3
"""Per-source-line summary of an ncu report captured with --set full --import-source on (-lineinfo build).

usage: source_hotspots.py REPORT.ncu-rep [top_n] > summary.txt
Reads `ncu -i REPORT --page source --print-source cuda,sass --csv` and adds up, per file:line, the executed
warp instructions, thread instructions and stall samples; then per file and per named region (line ranges of
warp_decode.cuh resolved from the function names in the file itself)."""
import csv, io, os, re, subprocess, sys
from collections import defaultdict

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 45
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                     capture_output=True, text=True).stdout
lines = defaultdict(lambda: [0, 0, 0, ""])          # (file, line) -> [warp inst, thread inst, samples, text]
cur = None
cols = None
for row in csv.reader(io.StringIO(out)):
    if not row:
        continue
    if row[0] == "File Path":
        cur = os.path.basename(row[1]); continue
    if row[0] == "Function Name":
        continue
    if row[0] == "Line No":
        cols = {name: i for i, name in enumerate(row)}
        continue
    if cur is None or cols is None or not row[0]:
        continue                                     # SASS rows hang under their source line
    try:
        key = (cur, int(row[0]))
        e = lines[key]
        e[0] += int(row[cols["Instructions Executed"]])
        e[1] += int(row[cols["Thread Instructions Executed"]])
        e[2] += int(row[cols["# Samples"]])
        e[3] = row[1].strip()
    except (ValueError, KeyError, IndexError):
        pass
tot_i = sum(e[0] for e in lines.values()) or 1
tot_t = sum(e[1] for e in lines.values()) or 1
tot_s = sum(e[2] for e in lines.values()) or 1
print(f"total warp-inst {tot_i}  thread-inst {tot_t}  (avg active lanes {tot_t / tot_i:.1f})  samples {tot_s}")
files = defaultdict(lambda: [0, 0])
for (f, _), e in lines.items():
    files[f][0] += e[0]; files[f][1] += e[2]
for f, (i, s) in sorted(files.items(), key=lambda kv: -kv[1][0]):
    print(f"{f:32s} inst {100 * i / tot_i:5.1f}%  samples {100 * s / tot_s:5.1f}%")
print("--- top lines by warp instructions")
for (f, ln), e in sorted(lines.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{f}:{ln:4d} inst {100 * e[0] / tot_i:5.1f}% lanes {e[1] / max(e[0], 1):4.1f} samp {100 * e[2] / tot_s:5.1f}%  {e[3][:100]}")
# regions of warp_decode.cuh: from each top-level function's line to the next one's
here = os.path.dirname(os.path.abspath(__file__))
src = os.path.join(here, "..", "..", "defl_b200", "csrc", "warp_decode.cuh")
if os.path.exists(src):
    marks = []
    for n, text in enumerate(open(src), 1):
        m = re.match(r"(?:template.*\n)?DEFL_DEV\s+[\w:<> ]+?\s+(\w+)\(", text)
        if m:
            marks.append((n, m.group(1)))
        elif text.startswith("struct "):
            marks.append((n, text.split()[1]))
    marks.append((10 ** 9, "end"))
    print("--- regions of warp_decode.cuh (function start to next function start)")
    for (a, name), (b, _) in zip(marks, marks[1:]):
        i = sum(e[0] for (f, ln), e in lines.items() if f == "warp_decode.cuh" and a <= ln < b)
        s = sum(e[2] for (f, ln), e in lines.items() if f == "warp_decode.cuh" and a <= ln < b)
        if i:
            print(f"{name:20s} lines {a:4d}-{min(b, 9999) - 1:4d} inst {100 * i / tot_i:5.1f}%  samples {100 * s / tot_s:5.1f}%")
