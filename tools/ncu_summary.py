#!/usr/bin/env python
"""Turns ncu output brought back in gpurun_out/ into the small text summaries committed under profiles/.

    python tools/ncu_summary.py launches gpurun_out/<tag>/launches.csv  > profiles/<name>_launches.txt
    python tools/ncu_summary.py full     gpurun_out/<tag>/prof.ncu-rep  > profiles/<name>_ncu_full.txt
    python tools/ncu_summary.py traffic  <rep> <plain log of the same command> [<rep2> <log2> ...] > profiles/ncu_traffic.json
"""
import collections
import csv
import io
import re
import subprocess
import sys

FULL_METRICS = [
    "gpu__time_duration.sum",
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__cycles_elapsed.max",
]


def launches(path):
    with open(path) as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    agg = collections.OrderedDict()
    order = []
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1e3 if u in ("ns", "nsecond") else v * 1e3 if u in ("ms", "msecond") else v
        name = re.sub(r"\(.*", "", row["Kernel Name"])
        name = re.sub(r"^void ", "", name)[:100]
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
        order.append((name, v))
    tot = sum(a[1] for a in agg.values())
    n = sum(a[0] for a in agg.values())
    print("# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare SHARES)")
    print("# %d launches, %.1f us total" % (n, tot))
    print("%12s %7s %6s  %s" % ("total_us", "share", "n", "kernel"))
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%12.1f %6.2f%% %6d  %s" % (a[1], 100 * a[1] / tot, a[0], k))
    own = sum(a[1] for k, a in agg.items() if k.startswith("cvpce::"))
    print("# own kernels (cvpce::*): %.1f us = %.2f%% of all GPU time in the profiled command" % (own, 100 * own / tot))


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    print("# ncu --set full --clock-control none: selected raw metrics per captured launch (%s)" % path)
    seen = collections.Counter()
    for d in data:
        name = re.sub(r"^void ", "", d[idx["Kernel Name"]])
        short = re.sub(r"\(.*", "", name)
        seen[short] += 1
        if seen[short] > 2:
            continue
        print("\n== %s   [launch #%d of this kernel]" % (name[:140], seen[short]))
        for m in FULL_METRICS:
            if m in idx:
                print("  %-78s %s %s" % (m, d[idx[m]], units[idx[m]]))
    print("\n# launches captured per kernel: " + ", ".join("%s x%d" % kv for kv in seen.items()))


STAGE_OF = {"region_pool_kernel": "region_pool_embed", "mac_nhwc_pixelsplit_kernel": "region_pool_embed", "mac_nhwc_stream_kernel": "region_pool_embed", "maxpool2x2_nhwc_kernel": "maxpool2x2", "crop_resize_sep_kernel": "crop_resize", "crop_resize_col_kernel": "crop_resize", "crop_resize_kernel": "crop_resize",
            "nms_batched_kernel": "nms", "select_candidates_kernel": "select_candidates", "rows_prepare_kernel": "rows_prepare",
            "sim_topk_tc_kernel": "sim_topk"}


def traffic(path, bench_log, *more):
    """profiles/ncu_traffic.json: DRAM bytes per launch (median over the captured launches) of each of our kernels, next to
    the algorithmic bytes per launch the profiled bench.py run itself reported (bench.py matches on those)."""
    import json
    import statistics
    # arguments: report, the plain log of the same command [, report2, log2 ...]; a stage is taken from the first pair that has it
    pairs = [(path, bench_log)] + [(more[i], more[i + 1]) for i in range(0, len(more) - 1, 2)]
    kernels_of = {}
    for pth, log in pairs:
        line = [ln for ln in open(log) if ln.startswith("{")][-1]
        kernels_of[pth] = json.loads(line)["kernels"]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    per = collections.defaultdict(list)
    src = {}
    for pth, _ in pairs:
      out = subprocess.run(["ncu", "-i", pth, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
      rows = list(csv.reader(io.StringIO(out)))
      hdr, units, data = rows[0], rows[1], rows[2:]
      idx = {h: i for i, h in enumerate(hdr)}
      for d in data:
        short = re.sub(r"[(<].*", "", re.sub(r"^void ", "", d[idx["Kernel Name"]])).replace("cvpce::", "")
        stage = STAGE_OF.get(short)
        if not stage or src.get(stage, pth) != pth:  # a stage is taken from the first report that has it
            continue
        b = 0.0
        for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            b += float(d[idx[m]].replace(",", "")) * scale[units[idx[m]]]
        per[stage].append(b)
        src[stage] = pth
    res = {}
    for stage, v in per.items():
        kernels = kernels_of[src[stage]]
        if stage in kernels and "algorithmic_bytes_per_launch" in kernels[stage]:
            res[stage] = {"dram_bytes_per_launch": statistics.mean(v), "launches_captured": len(v),
                          "algorithmic_bytes_per_launch": kernels[stage]["algorithmic_bytes_per_launch"],
                          "source": "ncu --set full --clock-control none, %s (mean over the captured launches)" % src[stage]}
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    {"launches": launches, "full": full, "traffic": traffic}[sys.argv[1]](*sys.argv[2:])
