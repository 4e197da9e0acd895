"""Derives profiles/r2_traffic.json (DRAM bytes per launch of the dominant kernel class, read by bench.py) and a
per-launch table from profiles/r2_ncu_step512.csv (ncu metrics of `python profiles/per_launch.py bf16 512`; the
LAST round of launches in the file is the profiled step)."""
import csv, json, os
here = os.path.dirname(os.path.abspath(__file__))
rows = [r for r in csv.reader(open(os.path.join(here, "r2_ncu_step512.csv"))) if len(r) > 10]
h = rows[0]
ki, mi, vi, ii = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("ID")
launches = {}
for r in rows[1:]:
    launches.setdefault(int(r[ii]), {"name": r[ki]})[r[mi]] = float(r[vi].replace(",", ""))
ids = sorted(launches)
# the profiled step = everything from the last gather_map launch on (project) plus the forward before it; keep the
# backward chain only: from the last rank1_pad_kernel to the end
last_rank1 = max(i for i in ids if "rank1_pad_kernel" in launches[i]["name"])
step = [launches[i] for i in ids if i >= last_rank1]
print("%-34s %9s %10s %10s %8s %7s" % ("kernel", "us", "rd MB", "wr MB", "tensor%", "L2hit%"))
conv = []
for l in step:
    nm = l["name"].split("(")[0].split("::")[-1][:34]
    print("%-34s %9.1f %10.1f %10.1f %8.1f %7.1f" % (nm, l["gpu__time_duration.sum"] / 1e3, l["dram__bytes_read.sum"] / 1e6,
          l["dram__bytes_write.sum"] / 1e6, l["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"], l["lts__t_sector_hit_rate.pct"]))
    if "deconv_" in l["name"]:
        conv.append(l["dram__bytes_read.sum"] + l["dram__bytes_write.sum"])
out = {"conv_tc": {"launches": len(conv), "dram_bytes_per_launch": sum(conv) / len(conv),
                   "source": "profiles/r2_ncu_step512.csv (ncu dram__bytes_read.sum + dram__bytes_write.sum, one config-2 step, 512 projections)"}}
json.dump(out, open(os.path.join(here, "r2_traffic.json"), "w"), indent=1)
print(json.dumps(out))
