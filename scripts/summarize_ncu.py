"""Summarise `ncu --set full` reports (one or more kernels each) into the text committed under profiles/ and
update profiles/traffic.json.  Usage: python scripts/summarize_ncu.py out.txt name=report.ncu-rep[:traffic_key] ..."""
import csv, io, json, os, subprocess, sys

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
           "launch__waves_per_multiprocessor", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
           "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "smsp__cycles_active.avg",
           "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
           "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "smsp__average_warp_latency_issue_stalled_wait.ratio",
           "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio"]
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def rows(rep):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(txt)))
    head, units, body = r[0], r[1], r[2:]
    return head, units, body


out_path, specs = sys.argv[1], sys.argv[2:]
tpath = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "traffic.json")
traffic = json.load(open(tpath)) if os.path.exists(tpath) else {}
lines = ["# ncu --set full --clock-control none (scripts/gpu_ncu_full.sh), summarised by scripts/summarize_ncu.py"]
for spec in specs:
    name, rest = spec.split("=", 1)
    rep, _, tkey = rest.partition(":")
    head, units, body = rows(rep)
    col = {h: i for i, h in enumerate(head)}
    total = 0.0
    for b in body:
        lines.append("== %s  (%s, grid %s, block %s)" % (name, b[col["Kernel Name"]][:90], b[col["Grid Size"]], b[col["Block Size"]]))
        for m in METRICS:
            if m in col:
                lines.append("  %-90s %s %s" % (m, b[col[m]], units[col[m]]))
        rd = float(b[col["dram__bytes_read.sum"]].replace(",", "")) * UNIT.get(units[col["dram__bytes_read.sum"]], 1.0)
        wr = float(b[col["dram__bytes_write.sum"]].replace(",", "")) * UNIT.get(units[col["dram__bytes_write.sum"]], 1.0)
        lines.append("  dram traffic of this launch (read+write): %.3f MB" % ((rd + wr) / 1e6))
        total += rd + wr
    if tkey:
        traffic[tkey] = total
        lines.append("  -> traffic.json[%s] = %.0f bytes (sum over the %d kernels of one call)" % (tkey, total, len(body)))
open(out_path, "w").write("\n".join(lines) + "\n")
json.dump(traffic, open(tpath, "w"), indent=1)
print("\n".join(lines))
