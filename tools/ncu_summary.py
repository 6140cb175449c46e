"""Summarise an .ncu-rep (raw metrics + hot SASS segments) into text for profiles/."""
import csv
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[0]
keys = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "smsp__inst_executed.sum", "sm__inst_executed.avg.per_cycle_elapsed",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "launch__shared_mem_per_block_dynamic"]
for r in rows[2:]:
    for k in keys:
        if k in hdr:
            print("%-85s %s %s" % (k, r[hdr.index(k)], rows[1][hdr.index(k)]))
    print("---")
sass = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True,
                      text=True).stdout
rows = list(csv.reader(sass.splitlines()))
hdr = rows[1]
ia, isrc, ie, it, isamp = (hdr.index(x) for x in ("Address", "Source", "Instructions Executed",
                                                  "Thread Instructions Executed", "# Samples"))
ins = []
for r in rows[2:]:
    if len(r) < len(hdr):
        if ins:
            break
        continue
    try:
        ins.append((r[ia], r[isrc].strip(), int(r[ie]), int(r[it]), int(r[isamp])))
    except ValueError:
        pass
tot = sum(x[2] for x in ins)
tots = sum(x[4] for x in ins)
print("SASS instructions %d, warp-instructions executed %d, samples %d" % (len(ins), tot, tots))
seg, cur = [], None
for k, (a, s, e, t, sm) in enumerate(ins):
    if cur is None or not (0.5 * cur["e0"] <= e <= 2 * cur["e0"]):
        if cur:
            seg.append(cur)
        cur = {"start": k, "e0": max(e, 1), "n": 0, "exec": 0, "thr": 0, "samp": 0, "first": s}
    cur["n"] += 1
    cur["exec"] += e
    cur["thr"] += t
    cur["samp"] += sm
    cur["end"] = k
seg.append(cur)
for sg in seg:
    if sg["exec"] / tot > 0.01 or sg["samp"] / tots > 0.01:
        print("%5d-%5d n=%4d exec=%5.1f%% samples=%5.1f%% lanes=%4.1f exec/inst=%d  %s" % (
            sg["start"], sg["end"], sg["n"], 100 * sg["exec"] / tot, 100 * sg["samp"] / tots,
            sg["thr"] / max(sg["exec"], 1), sg["exec"] // sg["n"], sg["first"][:48]))
if len(sys.argv) > 3:
    lo, hi = int(sys.argv[2]), int(sys.argv[3])
    for k in range(lo, hi):
        a, s, e, t, sm = ins[k]
        print(k, "%-64s" % s[:64], e // 1000000, round(t / max(e, 1), 1), sm)
