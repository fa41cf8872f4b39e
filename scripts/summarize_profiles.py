"""Turn gpurun_out/launches.csv and a full ncu report into the text summaries kept under profiles/."""
import collections
import csv
import subprocess
import sys

tag, launches, rep, cmd = sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4]
rows = [r for r in csv.reader(open(launches)) if len(r) > 10]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    name = r[ki].split("(")[0]
    agg[name][0] += 1
    agg[name][1] += v
tot = sum(v[1] for v in agg.values())
with open(f"profiles/{tag}_launches_summary.txt", "w") as f:
    f.write(f"ncu --metrics gpu__time_duration.sum --clock-control none ... {cmd}\n")
    f.write("durations in ns are cold-cache / serialised: compare SHARES with bench.py's `kernels` object\n")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"{v[1] / tot * 100:6.2f}%  n={v[0]:4d}  avg={v[1] / v[0] / 1e3:9.1f} us  {k}\n")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(raw.splitlines()))
h, units = rr[0], rr[1]
keep = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__grid_size", "launch__cluster_size",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max"]
with open(f"profiles/{tag}_ncu_full.txt", "w") as f:
    f.write(f"ncu --set full --clock-control none --import-source on ... {cmd}\n")
    for r in rr[2:]:
        for w in keep:
            if w in h:
                f.write(f"{w} = {r[h.index(w)]} {units[h.index(w)]}\n")
        f.write("---\n")
print(open(f"profiles/{tag}_launches_summary.txt").read())
print(open(f"profiles/{tag}_ncu_full.txt").read())
