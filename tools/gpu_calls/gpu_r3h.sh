#!/bin/bash
OUT=gpurun_out/r3h
mkdir -p $OUT
timeout 300 python tools/sim_ab.py > $OUT/plain.log 2>&1 &&
timeout 900 ncu --set full --import-source on --cache-control none --clock-control none -k 'regex:sim_select_kernel' -s 5 -c 1 -o $OUT/prof_sel python tools/sim_ab.py > $OUT/ncu.log 2>&1
echo "ncu exit $?"
ncu -i $OUT/prof_sel.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys,re
rows=list(csv.reader(sys.stdin))
hdr,units,vals=rows[0],rows[1],rows[2]
for h,u,v in zip(hdr,units,vals):
    if re.search(r'gpu__time_duration.sum|issue_stalled.*ratio|smsp__inst_executed.sum$|sm__warps_active.avg|l1tex__t_sector_hit_rate|lts__t_sector_hit_rate|dram__bytes_read.sum$|smsp__issue_active.avg.pct|launch__registers|launch__occupancy_limit|smsp__cycles_active.avg$|l1tex__throughput.avg.pct|lts__throughput.avg.pct|l1tex__m_xbar2l1tex_read_bytes.sum$|launch__waves|sm__cycles_elapsed.max$', h):
        print('%-100s %s %s' % (h,v,u))
" > $OUT/select_metrics.txt
ncu -i $OUT/prof_sel.ncu-rep --page source --csv --print-source sass 2>/dev/null > $OUT/select_sass.csv
ncu -i $OUT/prof_sel.ncu-rep --page source --csv --print-source cuda 2>/dev/null > $OUT/select_cuda.csv
rm -f $OUT/prof_sel.ncu-rep
python - <<'PY'
import csv
for name in ('cuda',):
    rows=list(csv.reader(open('gpurun_out/r3h/select_%s.csv'%name)))
    h=None
    for i,r in enumerate(rows[:5]):
        if '# Samples' in r: h=i
    if h is None: print('no header', rows[:2]); continue
    hd=rows[h]; si=hd.index('# Samples'); ie=hd.index('Instructions Executed'); so=hd.index('Source')
    stall_cols=[i for i,x in enumerate(hd) if x.startswith('stall_') and 'Not Issued' not in x]
    data=[]
    for r in rows[h+1:]:
        try: data.append((int(r[si] or 0), int(r[ie] or 0), r[so][:100], {hd[i]:int(r[i] or 0) for i in stall_cols}))
        except: pass
    tot=sum(d[0] for d in data)
    print('samples',tot)
    for d in sorted(data,key=lambda x:-x[0])[:22]:
        st=sorted(d[3].items(), key=lambda x:-x[1])[:3]
        print('%5.1f%% inst %9d  %-100s %s' % (100*d[0]/max(tot,1), d[1], d[2], st))
PY
cat $OUT/select_metrics.txt | grep -v "^LTS\|^SM_\|TriageCompute"
