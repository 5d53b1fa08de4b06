#!/bin/bash
# Round-2 ncu visit, part 2: --set full of the unwrap / chamfer / carrier-field kernels, the temperature FFT family and the blurs
# (separate bounded runs so that each group gets launches; summaries are produced on the box, the reports are not copied back).
set -u
mkdir -p gpurun_out
export VISTAF_GRAPHS=0
ARGS="--steps 1 --warmup 3 --frames 32 --distinct 8 --no-cpu --no-maps"
cap() {  # name regex skip count
  timeout 400 ncu --target-processes application-only --set full --clock-control none --kernel-name regex:"$2" --launch-skip $3 --launch-count $4 \
    -f -o gpurun_out/cap_$1 python bench.py $ARGS > gpurun_out/cap_$1.log 2>&1; echo "ncu $1 rc=$?"
  python tools/ncu_raw.py gpurun_out/cap_$1.ncu-rep > gpurun_out/r02_ncu_full_$1.txt 2>&1
  rm -f gpurun_out/cap_$1.ncu-rep gpurun_out/cap_$1.log
}
cap ftp_unwrap_chamfer_field 'k_uf_merge|k_uf_flatten|k_uf_init_rows|k_ftp_unwrap_apply|k_ftp_residues|k_chamfer2|k_ftp_field|k_ftp_rowdft|k_ftp_assemble' 60 40
cap temp_fft_features 'k_temp_fft_cols_peaks|k_temp_fft_rows|k_temp_ifft_rows|k_temp_split|k_temp_band|k_temp_features|k_temp_gray' 14 14
cap blur 'k_blur_cols|k_blur_rows' 60 24
du -sh gpurun_out
