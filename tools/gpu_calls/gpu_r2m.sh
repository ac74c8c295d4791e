#!/bin/bash
mkdir -p gpurun_out/r2m
timeout 1200 python -m pytest tests -q -m gpu > gpurun_out/r2m/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/r2m/pytest_gpu.log
tail -6 gpurun_out/r2m/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2m/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/r2m/smoke.log
