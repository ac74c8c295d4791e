#!/bin/bash
# L2 / fabric metrics of the tcgen05 search at the C3 and C5-shard shapes (what bounds the kernel), plus the split-count test
OUT=gpurun_out/r2v
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "sim or topk" > $OUT/pytest_sim.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest_sim.log
for shape in 6400,8192,1024,5 25000,125000,512,5; do
  tag=$(echo $shape | tr , _)
  CMD="python tools/kbench.py sim --iters 1 --shape $shape"
  $CMD > $OUT/plain_$tag.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none -k 'regex:sim_topk_tc_kernel' -s 3 -c 1 -o $OUT/prof_$tag $CMD > $OUT/ncu_$tag.log 2>&1
  echo "ncu $tag exit $?"
  ncu -i $OUT/prof_$tag.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin))
hdr,units,vals=rows[0],rows[1],rows[2]
import re
for h,u,v in zip(hdr,units,vals):
    if re.search(r'lts__|xbar|l1tex__m_|dram__bytes|gpu__time_duration|sm__pipe_tensor|sm__cycles_elapsed.max|l1tex__throughput|lts__throughput|sm__inst_executed.sum|smsp__inst_executed.sum|launch__grid_size|sm__throughput', h):
        print('%-90s %s %s' % (h,v,u))
" > $OUT/l2_metrics_$tag.txt
  rm -f $OUT/prof_$tag.ncu-rep
done
wc -l $OUT/l2_metrics_*.txt
