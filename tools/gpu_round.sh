#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench lines for the three workloads, ncu launch list.  Outputs -> gpurun_out/
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as G; G.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
for w in multimodal ftp temperature; do
  python bench.py --workload $w --steps 3 --warmup 3 ${BENCH_EXTRA:-} > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"
  tail -c 600 gpurun_out/bench_$w.json
done
if [ "${NCU:-1}" = "1" ]; then
  ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_multimodal.csv \
    python bench.py --workload multimodal --steps 1 --warmup 3 --frames 16 --no-cpu > gpurun_out/ncu_launch.log 2>&1; echo "ncu rc=$?"
fi
