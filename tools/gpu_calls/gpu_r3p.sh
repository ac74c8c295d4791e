#!/bin/bash
OUT=gpurun_out/r3p
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_guards.py -q -m gpu -x > $OUT/pytest_guards.log 2>&1; echo "guards exit $?"; tail -12 $OUT/pytest_guards.log
