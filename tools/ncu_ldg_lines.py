"""Reads `ncu --page source --csv --print-source cuda,sass` on stdin: for every global load/store SASS line prints the CUDA source
line it belongs to with its executed count and L2 sectors."""
import csv
import sys

rows = list(csv.reader(sys.stdin))
cur = None
hdr = None
for r in rows:
    if len(r) > 7 and r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or len(r) != len(hdr):
        continue
    if r[0] not in ("", "Line No") and r[2] == "-":
        cur = (r[0], r[1].strip()[:100])
        continue
    sass = r[3]
    if ("LDG" in sass or "STG" in sass) and cur:
        ie = r[hdr.index("Instructions Executed")]
        sec = r[hdr.index("L2 Theoretical Sectors Global")] if "L2 Theoretical Sectors Global" in hdr else "?"
        try:
            if int(ie) < 1000000: continue
        except ValueError:
            continue
        print(f"line {cur[0]}: {cur[1]}\n     {sass.strip()[:70]}  executed={ie} sectors={sec}")
