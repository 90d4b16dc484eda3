"""Key metrics of the launches in an ncu report (development aid).  usage: ncu_keys.py REPORT [kernel-substring]"""
import csv, subprocess, sys
rep = sys.argv[1]; sub = sys.argv[2] if len(sys.argv) > 2 else ""
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines())); hdr, units = rows[0], rows[1]
want = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "smsp__average_warp_latency_per_inst_issued.ratio", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__grid_size"]
for r in rows[2:]:
    d = dict(zip(hdr, r))
    if sub not in d["Kernel Name"]: continue
    print("==", d["Kernel Name"][:60])
    for k in want:
        if k in d: print(f"  {k:70s} {d[k]} {units[hdr.index(k)]}")
    st = sorted(((float(d[k].replace(',', '')), k) for k in hdr if k.startswith("smsp__average_warps_issue_stalled") and k.endswith("per_issue_active.ratio") and d[k]), reverse=True)
    for v, k in st[:9]: print(f"  stall {k.split('stalled_')[1].split('_per_issue')[0]:28s} {v:.2f}")
