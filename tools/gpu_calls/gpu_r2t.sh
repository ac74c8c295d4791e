#!/bin/bash
OUT=gpurun_out/r2t
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "mac or rmac" > $OUT/pytest_mac.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest_mac.log
V=$PWD/cvpce_b200/variants
run() { timeout 120 python tools/mac_ab.py "$@" >> $OUT/mac_ab.jsonl 2>> $OUT/mac_ab.err; }
run --tune mac_kernel=1
for c in 1 2 3 4 6; do run --tune mac_ctas_per_sm=$c; done
for c in 1 2; do CVPCE_B200_SO=$V/libmac_u16.so run --tune mac_ctas_per_sm=$c; done
for c in 2 4 8; do CVPCE_B200_SO=$V/libmac_u4.so run --tune mac_ctas_per_sm=$c; done
run --tune mac_kernel=1 --B 128
for c in 1 2 4; do run --tune mac_ctas_per_sm=$c --B 128; done
run --tune mac_kernel=1 --hw 32
for c in 2 4; do run --tune mac_ctas_per_sm=$c --hw 32; done
cat $OUT/mac_ab.jsonl
