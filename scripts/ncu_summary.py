"""Key numbers of one `ncu --set full` report as markdown: python scripts/ncu_summary.py report.ncu-rep [title] [launch index in the report]"""
import csv, subprocess, sys

rep = sys.argv[1]
title = sys.argv[2] if len(sys.argv) > 2 else rep
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(raw.splitlines()))
idx = int(sys.argv[3]) if len(sys.argv) > 3 else 0
h, units, v = r[0], r[1], r[2 + idx]
col = {k: i for i, k in enumerate(h)}
want = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "smsp__inst_executed.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct"]
print(f"## {title}\n\n| metric | value | unit |\n|---|---|---|")
for k in want:
    if k in col:
        print(f"| `{k}` | {v[col[k]]} | {units[col[k]]} |")
tensor = [k for k in h if "pipe_tensor" in k and k not in want]
for k in tensor[:6]:
    print(f"| `{k}` | {v[col[k]]} | {units[col[k]]} |")
print("\nWarp stall reasons (cycles per issued instruction):\n\n| reason | value |\n|---|---|")
st = [(k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), float(v[i] or 0))
      for i, k in enumerate(h) if "issue_stalled" in k and k.endswith("_per_issue_active.ratio")]
for k, x in sorted(st, key=lambda kv: -kv[1])[:8]:
    print(f"| {k} | {x:.2f} |")
