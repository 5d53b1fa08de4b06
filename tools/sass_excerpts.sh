#!/bin/bash
# cuobjdump -sass excerpts of the hot kernels of lib/libvistaf.so -> profiles/r02_sass_hot_kernels.txt
# (instruction mix per kernel + the first lines of its inner loop; the full dump is ~40 MB)
LIB=vistaf-roboskin-vision-integrated-multimodal-sensor_b200/lib/libvistaf.so
OUT=profiles/r02_sass_hot_kernels.txt
: > $OUT
for k in k_irls_accum k_sel_s1 k_blur_rows k_blur_cols k_temp_fft_cols_peaks k_temp_features k_ftp_field k_uf_merge k_chamfer2 k_ftp_unwrap_heap k_telea_march; do
  cuobjdump -sass -fun "$(cuobjdump -elf $LIB 2>/dev/null | grep -o "_Z[A-Za-z0-9_]*${k}[A-Za-z0-9_]*" | sort -u | head -1)" $LIB > /tmp/sass_$k.txt 2>/dev/null
  echo "==== $k : $(grep -c '^ *\/\*[0-9a-f]\{4\}\*\/' /tmp/sass_$k.txt) SASS instructions" >> $OUT
  grep -o '^ *\/\*[0-9a-f]\{4\}\*\/ *[A-Z0-9_.]*' /tmp/sass_$k.txt | awk '{print $2}' | sed 's/\..*//' | sort | uniq -c | sort -rn | head -14 >> $OUT
  grep -v '^ *\/\* 0x' /tmp/sass_$k.txt | grep -m1 -A16 'FFMA\|DFMA\|ATOMS' | cut -c1-100 >> $OUT
done
wc -l $OUT
