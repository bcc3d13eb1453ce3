"""Digest of an ncu launch list (--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv):
per-kernel time / share / DRAM bytes of the last complete training step, and the tc_gemm traffic constant that
bench.py quotes as roofline.traffic.   python tools/ncu_digest.py <launches.csv> <step.txt> [<tc_traffic.json>]"""
import collections
import csv
import json
import sys

src, out_txt = sys.argv[1], sys.argv[2]
rows = [r for r in csv.reader(open(src)) if len(r) > 10]
ix = {h: i for i, h in enumerate(rows[0])}
recs = {}
for r in rows[1:]:
    try:
        rid = int(r[ix["ID"]])
    except ValueError:
        continue
    d = recs.setdefault(rid, {"name": r[ix["Kernel Name"]]})
    d[r[ix["Metric Name"]]] = float(r[ix["Metric Value"]].replace(",", ""))
ids = sorted(recs)
starts = [k for k, i in enumerate(ids) if "obs_s4d" in recs[i]["name"]]   # first kernel of a step
a, b = starts[-2], starts[-1]
step = [recs[i] for i in ids[a:b]]
tot = sum(x["gpu__time_duration.sum"] for x in step)
agg = collections.OrderedDict()
for x in step:
    n = x["name"].split("(")[0]
    n = "tc::tc_gemm_kernel<*>" if "tc_gemm_kernel" in n else n
    g = agg.setdefault(n, [0, 0.0, 0.0, 0.0])
    g[0] += 1; g[1] += x["gpu__time_duration.sum"]; g[2] += x["dram__bytes_read.sum"]; g[3] += x["dram__bytes_write.sum"]
lines = [f"# {src}: last complete step = launches {a}..{b - 1} of {len(ids)} profiled",
         "# per-launch times are cold-cache and serialised by ncu: compare SHARES with bench.py's CUDA-event figures, not absolutes",
         f"# step total {tot / 1e6:.3f} ms over {len(step)} launches",
         f"{'kernel':62s} {'n':>3s} {'ms':>8s} {'share':>6s} {'dram_rd_MB':>10s} {'dram_wr_MB':>10s} {'GB/s':>7s}"]
for n, g in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    lines.append(f"{n[:62]:62s} {g[0]:3d} {g[1] / 1e6:8.3f} {g[1] / tot:6.3f} {g[2] / 1e6:10.1f} {g[3] / 1e6:10.1f} "
                 f"{(g[2] + g[3]) / g[1]:7.0f}")
lines.append("# every launch of the step, in order:")
for k, x in enumerate(step):
    lines.append(f"{k:3d} {x['name'].split('(')[0][:70]:70s} {x['gpu__time_duration.sum'] / 1e3:9.1f} us  "
                 f"rd {x['dram__bytes_read.sum'] / 1e6:8.1f} MB  wr {x['dram__bytes_write.sum'] / 1e6:8.1f} MB")
open(out_txt, "w").write("\n".join(lines) + "\n")
print("\n".join(lines[:20]))
if len(sys.argv) > 3:
    fam = [v for k, v in agg.items() if "tc_gemm_kernel" in k or "res_block_fwd_kernel" in k]   # the tcgen05 kernels
    g = [sum(v[i] for v in fam) for i in range(4)]
    json.dump({"dram_bytes_per_step": g[2] + g[3], "launches": g[0], "tc_share_of_step": g[1] / tot,
               "source": f"{src} (ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum "
                         "--clock-control none, python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e "
                         f"--no-store-feed --no-extra, last complete step = launches {a}..{b - 1})"},
              open(sys.argv[3], "w"), indent=1)
