#!/bin/bash
# Runs on the GPU box (via gpurun): parity tests, the default bench, the ncu launch list and one --set full capture
# of our kernels. Output under gpurun_out/<tag>/.
TAG=${1:-r1}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu.log
tail -3 $OUT/pytest_gpu.log
python bench.py --steps 3 --warmup 3 > $OUT/bench_c3.json 2> $OUT/bench_c3.err; echo "bench exit $?"
tail -c 600 $OUT/bench_c3.json
python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref exit $?"
PROF="python bench.py --from-head --shelves 4 --steps 1 --warmup 1 --no-cpu-baseline --no-extension-modes"
$PROF > $OUT/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $OUT/launches.csv $PROF > $OUT/ncu_launch.log 2>&1
echo "ncu launches exit $?"
$PROF > $OUT/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:select_candidates_kernel|nms_batched_kernel|crop_resize_sep_kernel|crop_resize_kernel|region_pool_kernel|mac_nhwc_pixelsplit_kernel|maxpool2x2_nhwc_kernel|rows_prepare_kernel|sim_topk_tc_kernel|sim_select_kernel' -c 40 -o $OUT/prof_own $PROF > $OUT/ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $OUT
