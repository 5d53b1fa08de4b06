#!/bin/bash
# Round-2 closing visit (one gpurun call, every step bounded): GPU tests, the three bench lines of the shipped build, the
# per-kernel launch lists at the bench's launch size, and --set full of the kernels changed last (column / row blur, the
# 16-byte selection sweeps, IRLS accumulate).  Summaries are made on the box; reports are not copied back.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r02_pytest_gpu_final.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r02_pytest_gpu_final.log
timeout 500 python bench.py --steps 5 --warmup 3 > gpurun_out/r02_bench_multimodal_final.json 2> gpurun_out/bench_mm.err; echo "bench mm rc=$?"
timeout 400 python bench.py --workload ftp --steps 5 --warmup 3 > gpurun_out/r02_bench_ftp_final.json 2> gpurun_out/bench_ftp.err; echo "bench ftp rc=$?"
timeout 400 python bench.py --workload temperature --steps 5 --warmup 3 > gpurun_out/r02_bench_temperature_final.json 2> gpurun_out/bench_temp.err; echo "bench temp rc=$?"
python - <<'PY'
import json
for n in ("multimodal", "ftp", "temperature"):
    try:
        d = json.loads(open(f"gpurun_out/r02_bench_{n}_final.json").read().strip().splitlines()[-1])
        print(n, round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "maps", d.get("e2e_maps", {}).get("value"), "roofline", round(d["roofline"]["frac"], 4),
              "parity", d.get("parity"), "cpu", d.get("cpu_baseline", {}).get("value"), d["clocks"])
    except Exception as e:
        print(n, "ERR", e)
PY
tools/gpu_launch_list.sh ftp > gpurun_out/ll_ftp.out 2>&1; tail -3 gpurun_out/ll_ftp.out | head -1
tools/gpu_launch_list.sh temperature > gpurun_out/ll_temp.out 2>&1
rm -f gpurun_out/ll_ftp.csv gpurun_out/ll_temperature.csv
export VISTAF_GRAPHS=0
ARGS="--steps 1 --warmup 3 --frames 32 --distinct 8 --no-cpu --no-maps"
cap() {  # name regex skip count
  timeout 400 ncu --target-processes application-only --set full --clock-control none --kernel-name regex:"$2" --launch-skip $3 --launch-count $4 \
    -f -o gpurun_out/cap_$1 python bench.py $ARGS > gpurun_out/cap_$1.log 2>&1; echo "ncu $1 rc=$?"
  python tools/ncu_raw.py gpurun_out/cap_$1.ncu-rep > gpurun_out/r02_ncu_full_$1.txt 2>&1
  rm -f gpurun_out/cap_$1.ncu-rep gpurun_out/cap_$1.log
}
if [ "${1:-all}" = "all" ]; then
cap blur_final 'k_blur_cols|k_blur_rows' 60 24
cap sel_irls_final 'k_sel_s1|k_sel_s2|k_irls_accum|k_irls_compact' 150 30
fi
cap temp_fft_final 'k_temp_fft_cols_peaks|k_temp_fft_rows|k_temp_ifft_rows' 12 9
du -sh gpurun_out
