"""Text summary of an .ncu-rep (read on the CPU box): headline metrics, pipe utilisation,
stall reasons per issue, and where the stall samples sit by execution count of the SASS line.

  python tools/ncu_summary.py gpurun_out/prof_quad_v6.ncu-rep > profiles/ncu_quad_v6_summary.txt
"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__cluster_size",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__shared_mem_per_block_static", "launch__waves_per_multiprocessor",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "lts__t_bytes.sum",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__cycles_active.avg",
        "sm__cycles_elapsed.avg", "sm__cycles_elapsed.avg.per_second"]


def run(args):
    return subprocess.run(["ncu", "-i", rep] + args, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                          text=True).stdout


rows = list(csv.reader(io.StringIO(run(["--page", "raw", "--csv"]))))
hdr, units = rows[0], rows[1]
for vals in rows[2:]:
    d = dict(zip(hdr, zip(units, vals)))
    print("# kernel:", d.get("Kernel Name", ("", "?"))[1])
    for k in KEYS:
        if k in d:
            print("%-75s %-16s %s" % (k, d[k][0], d[k][1]))
    for k in hdr:
        if k.startswith("sm__inst_executed_pipe_") and k.endswith(".avg.pct_of_peak_sustained_active"):
            v = d[k][1]
            if v and float(v) >= 1.0:
                print("%-75s %-16s %s" % (k, d[k][0], v))
    for k in hdr:
        if "issue_stalled" in k and k.endswith("per_issue_active.ratio"):
            v = d[k][1]
            if v and float(v) >= 0.02:
                print("%-75s %-16s %s" % (k, d[k][0], v))

src = list(csv.reader(io.StringIO(run(["--page", "source", "--csv", "--print-source", "sass"]))))
if len(src) > 2:
    h = src[1]
    ix = {n: i for i, n in enumerate(h)}
    data = [r for r in src[2:] if len(r) == len(h)]
    tot = sum(int(r[ix["# Samples"]] or 0) for r in data) or 1
    lv, cnt = collections.Counter(), collections.Counter()
    for r in data:
        e = int(r[ix["Instructions Executed"]] or 0)
        lv[e] += int(r[ix["# Samples"]] or 0)
        cnt[e] += 1
    print("\n# stall samples by execution count of the SASS line (hot loops share a count)")
    for e, s in sorted(lv.items(), key=lambda x: -x[1])[:8]:
        print("executed %10d x   %4d SASS lines   %6d samples  %5.1f%%" % (e, cnt[e], s, 100.0 * s / tot))
    print("\n# top SASS lines by stall samples")
    for r in sorted(data, key=lambda r: -int(r[ix["# Samples"]] or 0))[:25]:
        print("%6s samples  x%-9s %s" % (r[ix["# Samples"]], r[ix["Instructions Executed"]], r[ix["Source"]][:90]))
