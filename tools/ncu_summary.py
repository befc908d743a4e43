"""Analysis aid: one-screen summary of an ncu report (raw page) -> stdout.  python tools/ncu_summary.py rep [rep...]"""
import csv, io, subprocess, sys
KEYS = [
 "gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
 "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__thread_inst_executed.sum",
 "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__thread_inst_executed_per_inst_executed.pct",
 "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__instruction_throughput.avg.pct_of_peak_sustained_active",
 "sm__icc_request_hit_rate.pct", "gcc__cache_requests_type_instruction.sum.pct_of_peak_sustained_elapsed",
 "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
 "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
 "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_cbu.avg.pct_of_peak_sustained_active",
 "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_adu.avg.pct_of_peak_sustained_active",
 "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum", "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum",
 "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum",
 "dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
 "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
]
STALL = "smsp__average_warps_issue_stalled_"
for rep in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h, units = rows[0], rows[1]
    for v in rows[2:]:
        d = dict(zip(h, v)); u = dict(zip(h, units))
        print("==", rep, d.get("Kernel Name", ""), "grid", d.get("Grid Size", ""), "block", d.get("Block Size", ""))
        for k in KEYS:
            if k in d: print("  %-75s %s %s" % (k, d[k], u.get(k, "")))
        st = sorted(((float(d[k]), k) for k in d if k.startswith(STALL) and k.endswith("_per_issue_active.ratio") and d[k] not in ("", "n/a")), reverse=True)
        for val, k in st[:9]:
            print("  stall %-60s %.3f" % (k[len(STALL):-len("_per_issue_active.ratio")], val))
