#!/bin/bash
# Run on the GPU box (one GPU):  bash scripts/ncu_capture.sh <tag>
# 1. the bench command once without ncu (must exit 0), 2. launch list (gpu__time_duration per launch) of the same command,
# 3. ncu --set full of ONE whole steady-state training step (every kernel of the step, so that per-kernel DRAM traffic and the
#    algorithmic bytes bench.py reports refer to the same launches).  Raw outputs land in gpurun_out/; summarise them here with
#    python scripts/ncu_summarize.py gpurun_out/<tag>_launches.csv gpurun_out/<tag>_full_raw.csv <tag>
set -u
TAG=${1:-r02}
OUT=gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --repeats 1 --no-cpu-baseline --no-configs --recon-steps 64"
mkdir -p $OUT
$CMD > $OUT/${TAG}_plain.json 2> $OUT/${TAG}_plain.err || { echo "plain run failed"; tail -5 $OUT/${TAG}_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu1.log 2>&1
# first launch of the 4th training step = the 4th assemble_kernel launch (3 warm-up steps precede it)
read SKIP COUNT < <(python - <<PY
import csv
rows = [r for r in csv.DictReader(l for l in open("$OUT/${TAG}_launches.csv") if l.startswith('"'))]
idx = [i for i, r in enumerate(rows) if "assemble_kernel" in r["Kernel Name"]]
print(idx[3], idx[4] - idx[3])
PY
)
echo "one step = launches [$SKIP, $SKIP+$COUNT)"
ncu --set full --clock-control none -s $SKIP -c $COUNT -o /tmp/${TAG}_full -f $CMD > $OUT/${TAG}_ncu2.log 2>&1
# the report of a whole step is > 64 MiB (the limit of what gpurun brings back): export its raw page here, keep the report on the box
ncu -i /tmp/${TAG}_full.ncu-rep --page raw --csv > $OUT/${TAG}_full_raw.csv
# a small report WITH source correlation of the dominant kernel (two launches of conv3x3_persist<32>)
ncu --set full --clock-control none --import-source on -k regex:conv3x3_persist -s 30 -c 2 -o $OUT/${TAG}_top -f $CMD > $OUT/${TAG}_ncu3.log 2>&1
ls -la $OUT/${TAG}_full_raw.csv $OUT/${TAG}_top.ncu-rep
