#!/bin/bash
# what does the selection kernel wait for?  ncu --set full with the caches left as the search leaves them
OUT=gpurun_out/r3g
mkdir -p $OUT
timeout 300 python tools/sim_ab.py > $OUT/plain.log 2>&1 &&
timeout 900 ncu --set full --cache-control none --clock-control none -k 'regex:sim_select_kernel' -s 5 -c 1 -o $OUT/prof_sel python tools/sim_ab.py > $OUT/ncu.log 2>&1
echo "ncu exit $?"
ncu -i $OUT/prof_sel.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys,re
rows=list(csv.reader(sys.stdin))
hdr,units,vals=rows[0],rows[1],rows[2]
for h,u,v in zip(hdr,units,vals):
    if re.search(r'gpu__time_duration.sum|issue_stalled.*ratio|warp_issue_stalled.*_per_warp_active|smsp__inst_executed.sum$|sm__warps_active|achieved_occupancy|l1tex__t_sector_hit_rate|lts__t_sector_hit_rate|dram__bytes_read.sum$|smsp__issue_active.avg.pct|launch__registers|launch__occupancy_limit|smsp__cycles_active.avg$|l1tex__throughput.avg.pct|lts__throughput.avg.pct|l1tex__data_bank_conflicts|smsp__inst_executed_op_shared|sm__inst_executed_pipe_lsu|l1tex__m_xbar2l1tex_read_bytes.sum$|smsp__average_warp', h):
        print('%-100s %s %s' % (h,v,u))
" > $OUT/select_metrics.txt
ncu -i $OUT/prof_sel.ncu-rep --page source --csv 2>/dev/null > $OUT/select_source.csv
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/r3g/select_source.csv')))
print(rows[0][:12])
# find columns
hdr=rows[0]
def col(name):
    for i,h in enumerate(hdr):
        if h.strip()==name: return i
    return None
ci=col('# Samples (All)') or col('Samples (All)')
print('cols', [h for h in hdr][:30])
PY
wc -c $OUT/select_source.csv
rm -f $OUT/prof_sel.ncu-rep
cat $OUT/select_metrics.txt | head -80
