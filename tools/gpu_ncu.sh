#!/bin/bash
# ncu visit: launch list of one workload + one full capture of a named kernel.  Usage: WL=ftp KERN=k_telea_march SKIP=2 bash tools/gpu_ncu.sh
set -u
mkdir -p gpurun_out
WL=${WL:-multimodal}; KERN=${KERN:-k_telea_march}; SKIP=${SKIP:-2}; FR=${FR:-64}
python bench.py --workload $WL --steps 1 --warmup 3 --frames $FR --no-cpu > gpurun_out/ncu_plain.json 2> gpurun_out/ncu_plain.err; echo "plain rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c ${COUNT:-4000} --csv --log-file gpurun_out/launches_$WL.csv \
  python bench.py --workload $WL --steps 1 --warmup 3 --frames $FR --no-cpu > gpurun_out/ncu_launch.log 2>&1; echo "ncu list rc=$?"
if [ -n "$KERN" ]; then
ncu --set full --clock-control none --import-source on --kernel-name regex:$KERN --launch-skip $SKIP --launch-count 1 -f -o gpurun_out/full_$KERN \
  python bench.py --workload $WL --steps 1 --warmup 3 --frames $FR --no-cpu > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out/*.ncu-rep
fi
