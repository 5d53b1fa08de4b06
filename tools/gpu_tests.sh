#!/bin/bash
# quick GPU-box visit: parity tests + smoke + one short bench per workload (no ncu)
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q ${PYTEST_EXTRA:-} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as G; G.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
for w in ${WORKLOADS:-ftp temperature}; do
  python bench.py --workload $w --steps 3 --warmup 3 --no-cpu ${BENCH_EXTRA:-} > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_$w.json"))
    print("$w fps",round(d['value'],1),"e2e",round(d['e2e']['value'],1),"launches",d['gpu_launches'])
    for s in d['stages']: print(f"   {s['stage']:30s} {s['ms_per_frame']:.4f} ms/frame {100*s['share']:5.1f}%  launches {s['launches_per_step']}")
except Exception as e: print("bench parse failed", e); print(open("gpurun_out/bench_$w.err").read()[-2000:])
PY
done
