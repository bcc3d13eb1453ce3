"""Top SASS instructions by stall samples from `ncu -i X.ncu-rep --page source --csv` output.
Usage: python tools/ncu_top_stalls.py file.csv [N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = rows[1]
data = rows[2:]
isrc = hdr.index("Source")
isamp = hdr.index("# Samples")
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[isamp] or 0) for r in data)
print(rows[0][1], "total samples", tot)
order = sorted(range(len(data)), key=lambda k: -int(data[k][isamp] or 0))[:n]
for k in order:
    r = data[k]
    s = int(r[isamp] or 0)
    st = sorted([(int(r[i] or 0), hdr[i]) for i in stall], reverse=True)[:2]
    print(f"{k:5d} {s:6d} {100 * s / tot:5.1f}% {r[isrc][:100]:100s} {st}")
