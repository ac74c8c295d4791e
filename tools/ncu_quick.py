#!/usr/bin/env python
"""Quick look at one .ncu-rep: headline metrics, stall reasons, hottest SASS lines.  python tools/ncu_quick.py rep [n_hot]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
nhot = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'l1tex__t_sector_hit_rate.pct',
        'lts__t_sector_hit_rate.pct', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'launch__shared_mem_per_block_dynamic', 'smsp__warps_eligible.avg.per_cycle_active', 'sm__cycles_elapsed.max']
idx = {h: i for i, h in enumerate(hdr)}
for d in rows[2:]:
    print("==", d[idx['Kernel Name']][:120])
    for k in keys:
        if k in idx:
            print('  %-78s %s %s' % (k, d[idx[k]], units[idx[k]]))
    for i, h in enumerate(hdr):
        if 'issue_stalled' in h and 'ratio' in h and 'not_issued' not in h:
            try:
                v = float(d[i])
            except ValueError:
                continue
            if v > 0.3:
                print('  stall %-60s %.2f' % (h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''), v))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
# one table per profiled launch, each starting with its own header row; show the first launch of every kernel name
starts = [i for i, r in enumerate(rows) if 'Source' in r and 'Instructions Executed' in r]
names = [d[idx['Kernel Name']] for d in rows_raw[2:]] if (rows_raw := list(csv.reader(io.StringIO(raw)))) else []
seen = set()
for t, hi in enumerate(starts):
    name = names[t] if t < len(names) else "launch %d" % t
    if name in seen:
        continue
    seen.add(name)
    hdr = rows[hi]
    end = starts[t + 1] if t + 1 < len(starts) else len(rows)
    data = [r for r in rows[hi + 1:end] if len(r) == len(hdr)]
    ia, ins, smp = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('# Samples')
    ti = sum(int(r[ins]) for r in data if r[ins].isdigit()) or 1
    ts = sum(int(r[smp]) for r in data if r[smp].isdigit()) or 1
    print("-- %s: hottest SASS by stall samples (idx, %%inst, %%samples)" % name[:80])
    hot = sorted(range(len(data)), key=lambda n: -int(data[n][smp] if data[n][smp].isdigit() else 0))[:nhot]
    for n in sorted(hot):
        r = data[n]
        print('%5d %6.2f%%i %6.2f%%s  %s' % (n, 100 * int(r[ins] or 0) / ti, 100 * int(r[smp] or 0) / ts, r[ia].strip()[:100]))
