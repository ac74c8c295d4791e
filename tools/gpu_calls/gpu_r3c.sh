#!/bin/bash
# re-score with 16 candidates in flight: parity, per-kernel durations (ncu launch list), same-box A/B
OUT=gpurun_out/r3c
mkdir -p $OUT
rm -f $OUT/sim_ab.jsonl
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_guards.py tests/test_gpu_sharded.py tests/test_gpu_full_size.py -q -m gpu -k "sim or search or topk or sharded or hot_path or classifier" > $OUT/pytest_sim.log 2>&1; rc=$?; echo "pytest exit $rc"; tail -4 $OUT/pytest_sim.log
OLD=$PWD/cvpce_b200/variants/libold.so
for rep in 1 2; do
  CVPCE_B200_SO=$OLD timeout 300 python tools/sim_ab.py >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err
  timeout 300 python tools/sim_ab.py >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err
done
cat $OUT/sim_ab.jsonl
timeout 300 python tools/sim_ab.py > $OUT/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k 'regex:sim_topk_tc_kernel|sim_select_kernel' --csv --log-file $OUT/ncu.csv python tools/sim_ab.py > $OUT/ncu.log 2>&1
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open('gpurun_out/r3c/ncu.csv')) if len(r) > 10 and r[0].isdigit()]
by = collections.OrderedDict()
for r in rows:
    by.setdefault(r[0], {"k": r[4][:32], "grid": r[8]})[r[12]] = r[14]
seen = collections.Counter()
for i, d in by.items():
    key = (d["k"], d["grid"])
    seen[key] += 1
    if seen[key] in (3, 15, 27):
        print(d["k"], d["grid"], d.get("gpu__time_duration.sum"), d.get("smsp__inst_executed.sum"))
PY
rm -f $OUT/ncu.csv
