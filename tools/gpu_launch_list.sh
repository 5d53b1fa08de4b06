#!/bin/bash
# Per-kernel durations at the bench's launch size (64 frames per launch, one lane): tools/gpu_launch_list.sh ftp|temperature
set -u
W=${1:-ftp}
mkdir -p gpurun_out
export VISTAF_GRAPHS=0
ARGS="--workload $W --frames 64 --chunk 64 --chunk-temp 64 --lanes-ftp 1 --lanes-temp 1 --distinct 8 --steps 1 --warmup 3 --no-cpu --no-maps"
timeout 300 python bench.py $ARGS > gpurun_out/ll_${W}_plain.json 2> gpurun_out/ll_${W}_plain.err; echo "plain rc=$?"
timeout 500 ncu --target-processes application-only --metrics gpu__time_duration.sum --clock-control none -c 5000 --csv \
  --log-file gpurun_out/ll_${W}.csv python bench.py $ARGS > gpurun_out/ll_${W}.log 2>&1; echo "ncu list rc=$?"
python tools/ncu_summary.py gpurun_out/ll_${W}.csv full > gpurun_out/r02_${W}64_launches_summary.txt 2>&1
head -45 gpurun_out/r02_${W}64_launches_summary.txt
rm -f gpurun_out/ll_${W}.log
