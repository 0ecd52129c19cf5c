"""Key metrics of one `ncu --set full` capture: python scripts/ncu_summary.py rep.ncu-rep > profiles/x.txt"""
import csv, subprocess, sys
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct",
        "sm__pipe_tensor_cycles_active.avg.pct", "sm__pipe_tc_cycles_active.avg.pct", "sm__warps_active.avg.pct",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block",
        "sm__throughput.avg.pct", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct", "l1tex__throughput.avg.pct",
        "smsp__issue_active.avg.pct", "launch__occupancy_limit", "sm__cycles_elapsed.avg ", "lts__t_bytes.sum ",
        "smsp__inst_executed.sum ", "l1tex__data_bank_conflicts_pipe_lsu.sum "]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, u = rows[0], rows[1]
for r in rows[2:]:
    print("kernel:", r[h.index("Kernel Name")][:120])
    for i, n in enumerate(h):
        if any((n + " ").startswith(w) for w in WANT) and "per_second" not in n and ".pct_of_peak_sustained_elapsed" not in n[len("dram__bytes"):] if n.startswith("dram__bytes") else any((n + " ").startswith(w) for w in WANT):
            print(f"  {n:75s} {r[i]:>16s} {u[i]}")
