#!/usr/bin/env python
"""Summarise one kernel launch of an `ncu --set full` report (.ncu-rep, read here with `ncu -i ... --page raw --csv`):
writes the text summary kept under profiles/ and, with --json, the record bench.py quotes for `roofline.traffic` and
`roofline.executed_inst_per_key_ncu` (profiles/walk_kernel_ncu.json) together with the commit, date and command it came from.

  python tools/ncu_summary.py gpurun_out/ncu_walk_r02_a.ncu-rep --keys 77594624 --out profiles/ncu_walk_r02_a_summary.txt \
      --json profiles/walk_kernel_ncu.json --command "python bench.py --steps 1 --warmup 3 --windows-per-step 2 --no-cpu-baseline" --note "..."
"""
import argparse
import csv
import datetime
import io
import json
import subprocess
import sys

KEEP = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__cycles_active.sum", "sm__cycles_elapsed.avg.per_second",
    "sass__inst_executed_local_loads", "sass__inst_executed_local_stores", "sass__inst_executed_shared_loads", "sass__inst_executed_global_loads",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_imc_miss_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
]
SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}


def num(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--kernel", default=None, help="substring of the kernel name (default: the first launch in the report)")
    ap.add_argument("--index", type=int, default=0, help="which matching launch")
    ap.add_argument("--keys", type=int, default=0, help="keys one launch processes (for the per-key figures)")
    ap.add_argument("--out", default=None)
    ap.add_argument("--json", default=None)
    ap.add_argument("--command", default="")
    ap.add_argument("--note", default="")
    a = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", a.report, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ki = hdr.index("Kernel Name")
    sel = [r for r in data if a.kernel is None or a.kernel in r[ki]]
    if not sel:
        sys.exit("no launch of %r in %s" % (a.kernel, a.report))
    r = sel[min(a.index, len(sel) - 1)]
    col = {h: i for i, h in enumerate(hdr)}
    val = {h: (r[i], units[i]) for h, i in col.items()}
    try:
        commit = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    except OSError:
        commit = ""
    date = datetime.datetime.now(datetime.timezone.utc).strftime("%Y-%m-%dT%H:%M:%SZ")
    lines = ["# ncu --set full --clock-control none, %s; %s" % (r[ki], a.note or a.command),
             "# report %s, summarised %s at commit %s; command: %s" % (a.report, date, commit, a.command)]
    for h in ("Kernel Name", "Block Size", "Grid Size"):
        lines.append("%s [] = %s" % (h, r[col[h]]))
    for h in KEEP:
        if h in val:
            lines.append("%s [%s] = %s" % (h, val[h][1], val[h][0]))
    warp_inst = num(val["smsp__inst_executed.sum"][0])
    cyc = num(val["sm__cycles_active.sum"][0]) if "sm__cycles_active.sum" in val else None
    alu_pct = num(val["sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"][0])
    dram = sum(num(val[k][0]) * SCALE.get(val[k][1], 1.0) for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    dur_ms = num(val["gpu__time_duration.sum"][0]) * {"ms": 1.0, "us": 1e-3, "s": 1e3, "ns": 1e-6}.get(val["gpu__time_duration.sum"][1], 1.0)
    rec = {"kernel": r[ki], "source": a.out or a.report, "report": a.report, "commit": commit, "date": date, "command": a.command,
           "keys_per_launch": a.keys, "dram_bytes_per_launch": dram, "duration_ms_under_ncu": dur_ms}
    if a.keys:
        rec["thread_inst_per_key"] = warp_inst * 32 / a.keys
        lines.append("# executed instructions per key = %.1f (warp instructions x 32 / %d keys)" % (rec["thread_inst_per_key"], a.keys))
        lines.append("# DRAM bytes per key = %.2f" % (dram / a.keys))
        if cyc and alu_pct is not None:
            # the ALU pipe takes one warp instruction per 2 cycles per SM sub-partition = 2 per SM cycle
            rec["alu_pipe_inst_per_key"] = alu_pct / 100.0 * cyc * 2 * 32 / a.keys
            lines.append("# ALU-pipe instructions per key = %.1f (pipe_alu %% x active SM cycles x 2 warp-inst/cycle x 32 / keys)" % rec["alu_pipe_inst_per_key"])
    text = "\n".join(lines) + "\n"
    if a.out:
        with open(a.out, "w") as f:
            f.write(text)
    else:
        sys.stdout.write(text)
    if a.json:
        with open(a.json, "w") as f:
            json.dump(rec, f, indent=1)
            f.write("\n")


if __name__ == "__main__":
    main()
