"""Compact per-launch table from `ncu -i X.ncu-rep --page raw --csv` (one row per profiled launch).
Usage: python tools/ncu_full_table.py raw.csv out.csv "<header comment>" """
import csv
import sys

COLS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "launch__grid_size", "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active"]
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
kn = ix["Kernel Name"]
with open(sys.argv[2], "w") as f:
    f.write(f"# {sys.argv[3]}\n")
    f.write("# units: " + ", ".join(f"{c}={units[ix[c]]}" for c in COLS if c in ix) + "\n")
    f.write("launch,kernel," + ",".join(COLS) + "\n")
    for k, r in enumerate(data):
        name = r[kn].split("(")[0].replace("void ", "").replace("tc::", "").replace("(int)", "").replace("(unsigned int)", "")
        name = name.replace("rb::<unnamed>::", "")
        f.write(f'{k},"{name}",' + ",".join(r[ix[c]].replace(",", "") if c in ix else "" for c in COLS) + "\n")
print(open(sys.argv[2]).read()[:3000])
