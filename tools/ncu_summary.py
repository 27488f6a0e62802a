"""Condense an `ncu --set full` report into the 3-column (metric,unit,value) summaries kept under profiles/:
    ncu -i gpurun_out/x.ncu-rep --page raw --csv > /tmp/raw.csv
    python tools/ncu_summary.py /tmp/raw.csv "how it was captured" > profiles/rN_ncu_full_<what>.csv
One block per captured launch (kernel name, launch id), the metrics below only."""
import csv
import sys

KEEP = [
    "gpu__time_duration.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__m_xbar2l1tex_read_bytes.sum",
    "lts__t_sector_hit_rate.pct", "lts__t_sectors.sum",
    "launch__block_size", "launch__cluster_size", "launch__grid_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic",
    "sm__cycles_elapsed.avg", "sm__cycles_elapsed.avg.per_second",
    "sm__ops_path_tensor_op_utchmma_src_fp16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__inst_issued.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct",
    "l1tex__t_bytes_pipe_lsu_mem_global_op_st.sum",
    "l1tex__t_bytes_pipe_lsu_mem_global_op_ld.sum",
    "lts__t_sectors_op_atom.sum", "lts__t_sectors_op_red.sum",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
]


def main():
    path = sys.argv[1]
    how = sys.argv[2] if len(sys.argv) > 2 else ""
    with open(path, newline="") as f:
        rows = [r for r in csv.reader(f) if r]
    # ncu may print "==PROF==" lines first: the header is the row that starts with "ID"
    h = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    names, units, data = rows[h], rows[h + 1], rows[h + 2:]
    col = {n: i for i, n in enumerate(names)}
    print("metric,unit,value")
    for r in data:
        kname = r[col["Kernel Name"]].split("(")[0]
        print(f"kernel,,{kname} (launch id {r[col['ID']]}; ncu --set full --clock-control none; {how})")
        for m in KEEP:
            if m in col and r[col[m]] != "":
                print(f"{m},{units[col[m]]},{r[col[m]].replace(',', '')}")


if __name__ == "__main__":
    main()
