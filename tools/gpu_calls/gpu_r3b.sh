#!/bin/bash
# per-kernel durations (ncu launch list) of the search at the C3 shape: shipped kernel vs key lists vs keys off
OUT=gpurun_out/r3b
mkdir -p $OUT
OLD=$PWD/cvpce_b200/variants/libold.so
run() { # name, env so, tune...
  name=$1; so=$2; shift 2
  CVPCE_B200_SO=$so timeout 300 python tools/sim_ab.py "$@" > $OUT/plain_$name.log 2>&1 &&
  CVPCE_B200_SO=$so timeout 600 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum --clock-control none -k 'regex:sim_topk_tc_kernel|sim_select_kernel' --csv --log-file $OUT/ncu_$name.csv python tools/sim_ab.py "$@" > $OUT/ncu_$name.log 2>&1
  python - "$OUT/ncu_$name.csv" "$name" <<'PY'
import csv, sys, collections
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
by = collections.OrderedDict()
for r in rows:
    by.setdefault(r[0], {"k": r[4][:40], "grid": r[8]})[r[12]] = r[14]
out = []
for i, d in by.items():
    out.append((d["k"], d["grid"], d.get("gpu__time_duration.sum"), d.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"), d.get("smsp__inst_executed.sum")))
print("==", sys.argv[2])
for o in out[:60]:
    print(o)
PY
}
run old $OLD > $OUT/summary_old.txt
run keys "" > $OUT/summary_keys.txt
run nokeys "" --tune sim_keys=1 > $OUT/summary_nokeys.txt
rm -f $OUT/*.csv
head -40 $OUT/summary_old.txt; head -40 $OUT/summary_keys.txt
