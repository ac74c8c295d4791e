#!/bin/bash
# Round-2 final verification (1 GPU): full -m gpu suite, smoke(), the default bench line and kbench of every kernel with the final code.
OUT=gpurun_out/r2r
mkdir -p $OUT
python -m pytest tests -q -m gpu > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke exit $?"; tail -2 $OUT/smoke.log
python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "bench exit $?"
python tools/kbench.py nms select maxpool match crop pool prep sim > $OUT/kbench_all.jsonl 2> $OUT/kbench_all.err; echo "kbench exit $?"
cp gpurun_out/headline_mode_check.json $OUT/ 2>/dev/null
du -sh $OUT
