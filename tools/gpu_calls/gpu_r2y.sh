#!/bin/bash
# is the C3 search bound by what the fabric delivers into the SMs?  MMA-only modes of the linear-range experiment build
OUT=gpurun_out/r2y
mkdir -p $OUT
rm -f $OUT/sim_ab.jsonl
L=$PWD/cvpce_b200/variants/liblin.so
for t in "sim_plan=1 sim_epi_groups=1 sim_debug=0" "sim_plan=1 sim_epi_groups=1 sim_debug=1" "sim_plan=1 sim_epi_groups=1 sim_debug=2" "sim_plan=2 sim_epi_groups=1 sim_debug=0" "sim_plan=2 sim_epi_groups=1 sim_debug=1" "sim_plan=2 sim_epi_groups=1 sim_debug=2" "sim_plan=1 sim_epi_groups=1 sim_debug=2 sim_split=1" "sim_plan=1 sim_epi_groups=1 sim_debug=2 sim_split=3" "sim_plan=1 sim_epi_groups=1 sim_debug=0 sim_split=3"; do
  CVPCE_B200_SO=$L timeout 300 python tools/sim_ab.py --tune $t >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err
done
cat $OUT/sim_ab.jsonl
