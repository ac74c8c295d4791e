#!/bin/bash
# ncu --set full of the shipped warp-specialised R-MAC kernel (automatic choice) at B = 256 and B = 800, channels-last f32
OUT=gpurun_out/r5b
mkdir -p $OUT
for cfg in "256 nhwc f32" "800 nhwc f32"; do
  tag=$(echo $cfg | tr ' ' '_')
  timeout 120 python tools/rmac_one.py $cfg > $OUT/rmac_one_$tag.log 2>&1 &&
  timeout 300 ncu --set full --clock-control none --import-source on -k 'regex:rmac_staged' -s 3 -c 1 -o $OUT/prof_rmac_ws_$tag python tools/rmac_one.py $cfg > $OUT/ncu_rmac_ws_$tag.log 2>&1
  echo "ncu $tag exit $?"; cat $OUT/rmac_one_$tag.log | tail -3
  python tools/ncu_summary.py full $OUT/prof_rmac_ws_$tag.ncu-rep > $OUT/prof_rmac_ws_${tag}_ncu_full.txt 2>&1
done
rm -f $OUT/*.ncu-rep
head -14 $OUT/prof_rmac_ws_800_nhwc_f32_ncu_full.txt
