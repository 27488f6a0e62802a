"""ncu --csv launch list (long format, several metrics per launch) -> one line per launch:
    python tools/launch_table.py gpurun_out/x/launches_eval.csv "how it was captured" > profiles/r2_launches_ncx_eval.csv"""
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if r]
h = next(i for i, r in enumerate(rows) if r[0] == "ID")
col = {n: i for i, n in enumerate(rows[h])}
launches, order = {}, []
for r in rows[h + 1:]:
    key = r[col["ID"]]
    if key not in launches:
        launches[key] = {"kernel": r[col["Kernel Name"]].split("(")[0].replace("vqacx::", ""), "grid": r[col["Grid Size"]],
                         "block": r[col["Block Size"]]}
        order.append(key)
    launches[key][r[col["Metric Name"]]] = r[col["Metric Value"]].replace(",", "")
print(f"# {sys.argv[2] if len(sys.argv) > 2 else ''}")
print("id,kernel,grid,block,gpu__time_duration_us,dram_read_MB,dram_write_MB,tensor_pipe_active_pct_of_elapsed")
tot = 0.0
for k in order:
    L = launches[k]
    us = float(L.get("gpu__time_duration.sum", 0)) / 1e3
    tot += us
    print(f"{k},{L['kernel']},\"{L['grid']}\",\"{L['block']}\",{us:.2f},{float(L.get('dram__bytes_read.sum', 0)) / 1e6:.1f},"
          f"{float(L.get('dram__bytes_write.sum', 0)) / 1e6:.1f},"
          f"{float(L.get('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed', 0)):.1f}")
print(f"# sum of kernel durations: {tot:.1f} us over {len(order)} launches (serialised, cold-cache ncu pass)")
