"""Print the handful of ncu raw-page metrics that decide what bounds a kernel.  usage: python tools/ncu_keys.py report.ncu-rep"""
import csv, subprocess, sys, io
txt = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr = rows[0]
want = ["gpu__time_duration.sum", "smsp__issue_active.avg.pct", "sm__warps_active.avg.pct", "launch__registers_per_thread", "launch__occupancy_limit",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum", "smsp__average_warps_issue_stalled", "l1tex__t_sector_hit_rate.pct",
        "launch__grid_size", "sm__pipe_tensor_cycles_active", "sm__inst_executed_pipe_fma", "smsp__inst_executed_pipe_fma", "sm__pipe_fma_cycles_active"]
for r in rows[2:]:
    d = dict(zip(hdr, r))
    print("----", d["Kernel Name"][:70])
    for k in hdr:
        if any(w in k for w in want) and "per_second" not in k:
            try:
                if float(d[k]) == 0:
                    continue
            except ValueError:
                continue
            print("   %-95s %s" % (k, d[k]))
