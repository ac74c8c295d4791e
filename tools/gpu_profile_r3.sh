#!/bin/bash
# Final evidence run of round 2 (1 GPU): default bench, reference arm, ncu launch list of the step, ncu --set full of the own kernels on the
# default command (traffic per launch) and of the tcgen05 search at the C5 shard shape.  Output under gpurun_out/r3k/.
OUT=gpurun_out/r3k
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "bench exit $?"
python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_reference.json 2> $OUT/bench_reference.err; echo "ref exit $?"
PROF="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extension-modes --no-other-configs"
$PROF > $OUT/plain_prof.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file $OUT/launches.csv $PROF > $OUT/ncu_launch.log 2>&1
echo "ncu launches exit $?"
$PROF > $OUT/plain_prof2.log 2>&1 &&
ncu --set full --clock-control none -k 'regex:nms_batched_kernel|crop_resize_col_kernel|mac_nhwc_pixelsplit_kernel|mac_nhwc_stream_kernel|rows_prepare_kernel|sim_topk_tc_kernel|sim_select_kernel' -c 16 -o $OUT/prof_own $PROF > $OUT/ncu_full.log 2>&1
echo "ncu full exit $?"
$PROF > $OUT/plain_prof3.log 2>&1 &&
ncu --set full --clock-control none -k 'regex:maxpool2x2_nhwc_kernel' -c 8 -o $OUT/prof_maxpool $PROF > $OUT/ncu_full_maxpool.log 2>&1
echo "ncu maxpool exit $?"
C5="python tools/kbench.py sim --iters 1 --shape 100000,125000,512,5"
$C5 > $OUT/plain_c5.log 2>&1 &&
ncu --set full --clock-control none -k 'regex:sim_topk_tc_kernel|sim_select_kernel' -s 6 -c 4 -o $OUT/prof_sim_c5 $C5 > $OUT/ncu_full_c5.log 2>&1
echo "ncu c5 exit $?"
C3="python tools/kbench.py sim --iters 1 --shape 6400,8192,1024,5"
$C3 > $OUT/plain_c3.log 2>&1 &&
ncu --set full --clock-control none -k 'regex:sim_topk_tc_kernel|sim_select_kernel' -s 6 -c 4 -o $OUT/prof_sim_c3 $C3 > $OUT/ncu_full_c3.log 2>&1
echo "ncu c3 exit $?"
# the reports are too large to travel back (64 MiB limit): summarise them here, keep the text
for r in prof_own prof_maxpool prof_sim_c5 prof_sim_c3; do
  python tools/ncu_summary.py full $OUT/$r.ncu-rep > $OUT/${r}_ncu_full.txt 2>&1
  ncu -i $OUT/$r.ncu-rep --page raw --csv > $OUT/${r}_raw.csv 2>/dev/null
done
python tools/ncu_summary.py traffic $OUT/prof_own.ncu-rep $OUT/plain_prof2.log $OUT/prof_maxpool.ncu-rep $OUT/plain_prof3.log > $OUT/ncu_traffic.json 2> $OUT/ncu_traffic.err
python tools/ncu_summary.py launches $OUT/launches.csv > $OUT/launches.txt 2>&1
rm -f $OUT/*.ncu-rep $OUT/launches.csv
du -sh $OUT; ls -la $OUT
