"""Condense `ncu -i X.ncu-rep --page raw --csv` (one kernel) into the metric list kept under profiles/.
usage: python scripts/ncu_summary.py raw.csv > profiles/NAME.txt"""
import csv
import sys

KEEP = ["dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum",
        "launch__block_size", "launch__grid_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "sm__cycles_active.avg", "sm__cycles_elapsed.avg",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.per_cycle_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active"]


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    col = {h: i for i, h in enumerate(hdr)}
    print("kernel:", vals[col["Kernel Name"]][:160])
    print()
    for k in KEEP:
        if k in col:
            print("%s [%s] = %s" % (k, units[col[k]], vals[col[k]]))
    stalls = {}
    for h, i in col.items():
        if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued"):
            try:
                stalls[h[len("smsp__pcsamp_warps_issue_stalled_"):]] = float(vals[i].replace(",", ""))
            except ValueError:
                pass
    tot = sum(stalls.values())
    print()
    print("warp stall sampling, all samples = %d" % tot)
    for k, v in sorted(stalls.items(), key=lambda kv: -kv[1]):
        if v / tot >= 0.0005:
            print("  stall_%-24s %.1f%%" % (k, 100.0 * v / tot))


if __name__ == "__main__":
    main(sys.argv[1])
