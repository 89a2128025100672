"""ncu `--page raw --csv` (one wide row per launch) -> the `metric,unit,value` summaries kept under profiles/.
    python tools/ncu_summary.py raw.csv [kernel-substring] [launch-index]   (prints to stdout)"""
import csv
import sys

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__block_size", "launch__grid_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__cluster_size", "sm__cycles_elapsed.avg", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "l1tex__m_xbar2l1tex_read_bytes.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "smsp__sass_thread_inst_executed_op_fp64_pred_on.sum", "sm__sass_inst_executed_op_shared_ld.sum", "sm__sass_inst_executed_op_shared_st.sum"]


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    head, units, data = rows[0], rows[1], rows[2:]
    want = sys.argv[2] if len(sys.argv) > 2 else ""
    which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    kcol = head.index("Kernel Name")
    sel = [r for r in data if want in r[kcol]]
    if not sel:
        print("no launch matches", want, "| kernels:", sorted({r[kcol][:60] for r in data}))
        return
    r = sel[min(which, len(sel) - 1)]
    print("# kernel:", r[kcol][:140], "| launch %d of %d matching" % (min(which, len(sel) - 1), len(sel)))
    print("metric,unit,value")
    for name in KEEP:
        for i, h in enumerate(head):
            if h == name or h.endswith("." + name):
                print("%s,%s,%s" % (name, units[i], r[i].replace(",", "")))
                break


if __name__ == "__main__":
    main()
