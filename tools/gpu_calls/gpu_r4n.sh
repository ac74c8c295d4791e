#!/bin/bash
# staged R-MAC kernel: what the row-segment end of a channels-last map is made of (cycle counters)
OUT=gpurun_out/r4n
mkdir -p $OUT
CVPCE_B200_SO=cvpce_b200/variants/librmacprof.so timeout 200 python tools/rmac_prof.py > $OUT/rmac_prof.txt 2>&1; echo "prof exit $?"; grep -A11 "B=128 nhwc" $OUT/rmac_prof.txt
