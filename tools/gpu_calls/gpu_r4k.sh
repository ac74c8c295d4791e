#!/bin/bash
# evidence run with the staged R-MAC kernel in the tree: whole -m gpu suite, default bench line (now with extension_modes.rmac_l3),
# reference arm, ncu --set full of rmac_staged_kernel at B = 256 channels-last f32 / bf16 and NCHW f32
OUT=gpurun_out/r4k
mkdir -p $OUT
timeout 900 python -m pytest tests -x -q -m gpu > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/pytest_gpu.log
timeout 900 python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "bench exit $?"; tail -c 1500 $OUT/bench_default.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_reference.json 2> $OUT/bench_reference.err; echo "ref exit $?"
for cfg in "256 nhwc f32" "256 nhwc bf16" "256 nchw f32"; do
  tag=$(echo $cfg | tr ' ' '_')
  timeout 120 python tools/rmac_one.py $cfg > $OUT/rmac_one_$tag.log 2>&1 &&
  timeout 300 ncu --set full --clock-control none --import-source on -k 'regex:rmac_staged_kernel' -s 3 -c 1 -o $OUT/prof_rmac_$tag python tools/rmac_one.py $cfg > $OUT/ncu_rmac_$tag.log 2>&1
  echo "ncu $tag exit $?"
  python tools/ncu_summary.py full $OUT/prof_rmac_$tag.ncu-rep > $OUT/prof_rmac_${tag}_ncu_full.txt 2>&1
  ncu -i $OUT/prof_rmac_$tag.ncu-rep --page raw --csv > $OUT/prof_rmac_${tag}_raw.csv 2>/dev/null
done
rm -f $OUT/*.ncu-rep
du -sh $OUT
