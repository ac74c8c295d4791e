#!/bin/bash
# where the staged R-MAC kernel's block 0 spends its cycles (variant built with -DCVPCE_RMAC_PROF)
OUT=gpurun_out/r4d
mkdir -p $OUT
CVPCE_B200_SO=cvpce_b200/variants/librmacprof.so timeout 300 python tools/rmac_prof.py > $OUT/rmac_prof.txt 2>&1; echo "prof exit $?"; cat $OUT/rmac_prof.txt
