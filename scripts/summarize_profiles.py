"""Turn the ncu outputs of a gpurun call into the text summaries kept under profiles/ and into profiles/ncu_traffic.json
(per-launch dram bytes of the dominant kernels, read by bench.py for `roofline.traffic`).

    python scripts/summarize_profiles.py TAG launches.csv full_raw.csv "the profiled command" [extra_raw.csv:label ...]

launches.csv : ncu --metrics gpu__time_duration.sum --csv --log-file ...      (every launch with its device time)
*_raw.csv    : ncu -i report.ncu-rep --page raw --csv                          (one row per profiled launch, --set full)
"""
import collections
import csv
import json
import os
import sys

tag, launches, full_raw, cmd = sys.argv[1:5]
extras = sys.argv[5:]

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__cluster_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.max"]


def short(name):
    return name.split("(")[0].replace("void ", "").replace("ddc::", "")


def read_raw(path):
    rr = list(csv.reader(open(path)))
    h, units = rr[0], rr[1]
    ki = h.index("Kernel Name")
    rows = []
    for r in rr[2:]:
        d = {"kernel": short(r[ki])}
        for k in KEEP:
            if k in h:
                try:
                    d[k] = float(r[h.index(k)].replace(",", ""))
                except ValueError:
                    d[k] = r[h.index(k)]
                d[k + "#unit"] = units[h.index(k)]
        rows.append(d)
    return rows


def to_bytes(v, unit):
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


# ---- launch list
rows = [r for r in csv.reader(open(launches)) if len(r) > 10]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    agg[short(r[ki])][0] += 1
    agg[short(r[ki])][1] += v
tot = sum(v[1] for v in agg.values())
os.makedirs("profiles", exist_ok=True)
with open(f"profiles/{tag}_launches_summary.txt", "w") as f:
    f.write(f"ncu --metrics gpu__time_duration.sum --clock-control none ... {cmd}\n")
    f.write("durations are cold-cache / serialised under the profiler: compare SHARES with bench.py's `kernels` object\n")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"{v[1] / tot * 100:6.2f}%  n={v[0]:4d}  avg={v[1] / v[0] / 1e3:9.1f} us  {k}\n")

# ---- full sections
def dump(rows, f):
    for d in rows:
        f.write(f"[{d['kernel']}]\n")
        for k in KEEP:
            if k in d:
                f.write(f"  {k} = {d[k]} {d[k + '#unit']}\n")

full = read_raw(full_raw)
with open(f"profiles/{tag}_ncu_full.txt", "w") as f:
    f.write(f"ncu --set full --clock-control none ... {cmd}   (one denoise step: every launch once, in launch order)\n")
    dump(full, f)
    for e in extras:
        path, label = e.split(":", 1)
        f.write(f"\n==== {label}\n")
        dump(read_raw(path), f)

# ---- traffic of the kernel classes bench.py can name as dominant (S1, 1024 images per GPU): mean over the launches seen
cls_of = {"block_tail_kernel": "block_tail", "qkv_attn_tc_kernel": "qkv_attention", "attention_stream_kernel": "attention", "attention_fa_kernel": "attention",
          "attention_head_kernel": "attention"}
traffic = collections.defaultdict(list)
for d in full:
    base = d["kernel"].split("<")[0]
    cls = cls_of.get(base)
    if base == "gemm2_kernel" and d["kernel"].startswith("gemm2_kernel<0"):
        cls = "qkv"
    if cls and "dram__bytes_read.sum" in d:
        traffic[cls].append(to_bytes(d["dram__bytes_read.sum"], d["dram__bytes_read.sum#unit"]) +
                            to_bytes(d["dram__bytes_write.sum"], d["dram__bytes_write.sum#unit"]))
out = {"s1_1024": {c: {"bytes": sum(v) / len(v), "launches": len(v), "source": f"profiles/{tag}_ncu_full.txt"} for c, v in traffic.items()}}
with open("profiles/ncu_traffic.json", "w") as f:
    json.dump(out, f, indent=1, sort_keys=True)
print(open(f"profiles/{tag}_launches_summary.txt").read())
print(json.dumps(out, indent=1))
