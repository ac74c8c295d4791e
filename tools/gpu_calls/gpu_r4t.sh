#!/bin/bash
# staged R-MAC kernel: timeline of block 0 of the two staged kernels at B = 64 .. 512
OUT=gpurun_out/r4t
mkdir -p $OUT
CVPCE_B200_SO=cvpce_b200/variants/librmacprof.so timeout 200 python tools/rmac_prof.py > $OUT/rmac_prof.txt 2>&1; echo "prof exit $?"; cat $OUT/rmac_prof.txt
