"""Reads `ncu --page source --csv --print-source sass` on stdin and prints the SASS lines with the most L2 sectors / instructions."""
import csv
import sys

rows = list(csv.reader(sys.stdin))
hdr = next(r for r in rows if "Source" in r and "Address" in r)
ix = {n: hdr.index(n) for n in hdr}
want = [c for c in ("L2 Theoretical Sectors Global", "L2 Theoretical Sectors Global Ideal", "Instructions Executed", "Access Operation", "Access Size") if c in ix]
body = [r for r in rows if len(r) == len(hdr) and r is not hdr and r[0].startswith("0x")]
key = "L2 Theoretical Sectors Global"
def num(s):
    try: return int(s)
    except ValueError: return 0
body.sort(key=lambda r: -num(r[ix[key]]))
for r in body[:14]:
    print(" | ".join(f"{c}={r[ix[c]]}" for c in want), "|", r[ix["Source"]].strip()[:90])
