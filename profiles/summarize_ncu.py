"""Summary of an `ncu --set full` report: per kernel (first captured launch of each) duration, registers, occupancy, issue
rate, pipe instruction counts (FMA / ALU / XU / LSU / uniform), shared-memory wavefronts, L2 RED / ATOM sectors, DRAM bytes
and the top warp-stall reasons.  usage: python profiles/summarize_ncu.py gpurun_out/prof_X.ncu-rep > profiles/X_ncu_summary.txt"""
import csv
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
want = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__thread_inst_executed.sum",
        "sm__inst_executed.avg.per_cycle_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "smsp__inst_executed_pipe_fma.sum", "smsp__inst_executed_pipe_fmaheavy.sum", "smsp__inst_executed_pipe_fmalite.sum",
        "smsp__inst_executed_pipe_alu.sum", "smsp__inst_executed_pipe_xu.sum", "smsp__inst_executed_pipe_lsu.sum",
        "smsp__inst_executed_pipe_uniform.sum", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__inst_executed_op_shared_atom.sum", "lts__t_sectors_op_red.sum",
        "lts__t_sectors_op_atom.sum", "smsp__warps_eligible.avg.per_cycle_active", "smsp__warps_active.avg.per_cycle_active"]
stall = [h for h in hdr if "issue_stalled" in h and h.endswith("_per_issue_active.ratio") and "not_issued" not in h]
seen = set()
print(f"# {rep}: first captured launch of each kernel (ncu replays: durations are cold-cache, compare shares)")
for r in rows[2:]:
    name = r[idx["Kernel Name"]]
    key = name[:60]
    if key in seen:
        continue
    seen.add(key)
    print(f"\n== {name[:110]}")
    for w in want:
        if w in idx and r[idx[w]] != "":
            print(f"   {w:72s} {r[idx[w]]:>18s} {units[idx[w]]}")
    vals = sorted([(float(r[idx[h]].replace(",", "")), h.split("issue_stalled_")[1].replace("_per_issue_active.ratio", ""))
                   for h in stall if r[idx[h]]], reverse=True)[:7]
    print("   warp stalls per issued instruction: " + ", ".join(f"{b} {a:.2f}" for a, b in vals))
