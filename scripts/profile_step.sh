#!/bin/bash
# One GPU call that produces the raw material for profiles/<tag>_*: the launch list of a short bench run and one
# `ncu --set full` capture of a whole denoise step.  Run under gpurun; summarise afterwards (CPU box) with
#   ncu -i gpurun_out/<tag>_full.ncu-rep --page raw --csv > gpurun_out/<tag>_full_raw.csv
#   python scripts/summarize_profiles.py <tag> gpurun_out/<tag>_launches.csv gpurun_out/<tag>_full_raw.csv "<cmd>"
TAG=${1:-r02b}
CMD="python bench.py --steps 1 --warmup 3 --lean --no-cpu-baseline --k 500"
set -x
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || exit 1          # the profiled command must pass without the profiler first
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
# one whole step = the launches from the last-but-one unpatchify + DDIM launch to the next one (patchify runs once per loop)
SKIP=$(python - <<PY
import csv
rows = [r for r in csv.reader(open("gpurun_out/${TAG}_launches.csv")) if len(r) > 10]
ki = rows[0].index("Kernel Name")
names = [r[ki] for r in rows[1:]]
idx = [i for i, n in enumerate(names) if "head_ddim" in n]
print(idx[-2] + 1, min(idx[-1] - idx[-2], 40))
PY
)
set -- $SKIP
ncu --set full --clock-control none --import-source on -s $1 -c $2 -o gpurun_out/${TAG}_full -f $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
ls -la gpurun_out/${TAG}_full.ncu-rep
