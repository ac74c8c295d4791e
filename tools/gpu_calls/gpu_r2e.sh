#!/bin/bash
# full validation of the state after the host-side work: gpu tests, smoke, default bench
mkdir -p gpurun_out/r2e
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r2e/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/r2e/pytest_gpu.log
tail -5 gpurun_out/r2e/pytest_gpu.log
cp gpurun_out/headline_mode_check.json gpurun_out/r2e/ 2>/dev/null
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2e/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/r2e/smoke.log
timeout 900 python bench.py > gpurun_out/r2e/bench_default.json 2> gpurun_out/r2e/bench_default.err; echo "bench exit $?"
tail -c 3000 gpurun_out/r2e/bench_default.json; tail -5 gpurun_out/r2e/bench_default.err
