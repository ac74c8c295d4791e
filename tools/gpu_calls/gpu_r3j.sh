#!/bin/bash
# allocation-free MAC entry point: its test alone in a fresh process, then the whole -m gpu suite (it runs last)
OUT=gpurun_out/r3j
mkdir -p $OUT
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > $OUT/pytest_alone.log 2>&1; echo "alone exit $?"; tail -15 $OUT/pytest_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > $OUT/pytest_all.log 2>&1; echo "all exit $?"; tail -5 $OUT/pytest_all.log
