"""Per-kernel raw-page metrics of an .ncu-rep as a compact CSV + profiles/traffic.json.
usage: ncu_summary.py <rep> <out.csv> [traffic.json]"""
import csv, json, subprocess, sys
M = ("gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,launch__registers_per_thread,launch__occupancy_limit_registers,"
     "launch__occupancy_limit_shared_mem,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,"
     "smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,"
     "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,"
     "lts__t_sector_hit_rate.pct,smsp__warps_eligible.avg.per_cycle_active,"
     + ",".join(f"smsp__average_warps_issue_stalled_{k}_per_issue_active.ratio" for k in
                ("long_scoreboard", "short_scoreboard", "wait", "barrier", "math_pipe_throttle", "no_instruction", "lg_throttle",
                 "branch_resolving", "dispatch_stall", "mio_throttle")))
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv", "--metrics", M], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[0]
keep = [i for i, h in enumerate(hdr) if h in ("Kernel Name", "Grid Size", "Block Size") or "__" in h]
with open(sys.argv[2], "w", newline="") as f:
    w = csv.writer(f)
    for r in rows:
        rr = [r[i] for i in keep]
        rr[0] = rr[0][:40]
        w.writerow(rr)
if len(sys.argv) > 3:
    ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    units = {"Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Gbyte": 1e9}
    sr, sw = units.get(rows[1][ir], 1e6), units.get(rows[1][iw], 1e6)   # each column carries its own unit
    agg = {}
    for r in rows[2:]:
        k = r[hdr.index("Kernel Name")].split("(")[0].replace("void ", "").strip()
        agg.setdefault(k, []).append(float(r[ir]) * sr + float(r[iw]) * sw)
    out = {k: sum(v) / len(v) for k, v in agg.items()}
    out["k_solve"] = sum(v for k, v in out.items() if k.startswith("k_solve") or k == "k_classify")
    out["k_collide"] = out.get("k_broad", 0) + out.get("k_narrow", 0)
    out["_source"] = f"{sys.argv[2]}: dram__bytes_read.sum + dram__bytes_write.sum per launch (ncu --set full, cold L2), 4096 envs"
    json.dump(out, open(sys.argv[3], "w"), indent=1)
    print(out)
