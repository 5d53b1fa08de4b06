#!/bin/bash
# Round-2 ncu visit (one gpurun call, every step bounded by `timeout`):
#   1. plain run (must exit 0 before anything is profiled)
#   2. launch list of the first ~2 steps, eager launches (VISTAF_GRAPHS=0: one kernel per launch, as the graphs replay them)
#   3. --set full of the kernels DESIGN.md / VERDICT name (unwrap, chamfer, carrier field, temperature FFT family, features,
#      IRLS, blur, selection), a few launches each
set -u
mkdir -p gpurun_out
export VISTAF_GRAPHS=0
ARGS="--steps 1 --warmup 3 --frames 32 --distinct 8 --no-cpu --no-maps"
timeout 300 python bench.py $ARGS > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err; echo "plain rc=$?"
timeout 600 ncu --target-processes application-only --metrics gpu__time_duration.sum --clock-control none -c 3600 --csv \
  --log-file gpurun_out/r02_multimodal_launches.csv python bench.py $ARGS > gpurun_out/prof_list.log 2>&1; echo "ncu list rc=$?"
wc -l gpurun_out/r02_multimodal_launches.csv
K='k_uf_merge|k_uf_flatten|k_ftp_unwrap_apply|k_ftp_unwrap_heap|k_chamfer2|k_ftp_field|k_ftp_rowdft|k_temp_fft_cols_peaks|k_temp_fft_rows|k_temp_ifft_rows|k_temp_split|k_temp_band|k_temp_features|k_irls_accum|k_irls_compact|k_blur_cols|k_blur_rows|k_sel_s1|k_sel_s2|k_morph|k_ftp_assemble|k_t12_forward'
timeout 700 ncu --target-processes application-only --set full --clock-control none --kernel-name regex:"$K" --launch-skip 150 --launch-count 44 \
  -f -o gpurun_out/r02_full python bench.py $ARGS > gpurun_out/prof_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out/*.ncu-rep 2>/dev/null
# summaries are made on the box (the report itself may exceed what gpurun copies back)
python tools/ncu_summary.py gpurun_out/r02_multimodal_launches.csv > gpurun_out/r02_multimodal_launches_summary.txt 2>&1
python tools/ncu_raw.py gpurun_out/r02_full.ncu-rep > gpurun_out/r02_ncu_full_kernels.txt 2>&1
SZ=$(stat -c %s gpurun_out/r02_full.ncu-rep 2>/dev/null || echo 0)
if [ "$SZ" -gt 40000000 ]; then rm -f gpurun_out/r02_full.ncu-rep; fi
rm -f gpurun_out/prof_list.log gpurun_out/prof_full.log
du -sh gpurun_out
