#!/bin/bash
# staged R-MAC kernel: ring stages of 16 KB (shipped) against 32 KB (variant build), same box
OUT=gpurun_out/r4j
mkdir -p $OUT
timeout 150 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "rmac" > $OUT/pytest_rmac.log 2>&1; echo "rmac tests exit $?"; tail -3 $OUT/pytest_rmac.log
CVPCE_B200_SO=cvpce_b200/variants/librmac32k.so timeout 150 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "rmac" > $OUT/pytest_rmac_32k.log 2>&1; echo "rmac tests (32 KB stages) exit $?"; tail -3 $OUT/pytest_rmac_32k.log
for i in 1 2; do
timeout 200 python tools/kbench.py rmac --iters 7 > $OUT/kbench_rmac_16k_$i.jsonl 2> $OUT/kbench_rmac.err; echo "== 16 KB stages (run $i)"; grep "staged kernel\"" $OUT/kbench_rmac_16k_$i.jsonl | grep -v "B=32\|B=64"
CVPCE_B200_SO=cvpce_b200/variants/librmac32k.so timeout 200 python tools/kbench.py rmac --iters 7 > $OUT/kbench_rmac_32k_$i.jsonl 2> $OUT/kbench_rmac.err; echo "== 32 KB stages (run $i)"; grep "staged kernel\"" $OUT/kbench_rmac_32k_$i.jsonl | grep -v "B=32\|B=64"
done
