"""Developer tool: print the numbers of an evidence round (gpurun_out/<tag>_*) and copy the files to profiles/.
usage: round_summary.py <tag>"""
import csv, json, os, shutil, subprocess, sys
tag = sys.argv[1]
root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
go, pr = os.path.join(root, "gpurun_out"), os.path.join(root, "profiles")
print(open(f"{go}/{tag}_pytest_gpu.log").read().strip().splitlines()[-2:], open(f"{go}/{tag}_smoke.log").read().strip().splitlines()[-1])
d = json.load(open(f"{go}/{tag}_bench.json")); e = json.load(open(f"{go}/{tag}_bench20.json")); r = json.load(open(f"{go}/{tag}_bench_ref.json"))
print("200 steps:", d["value"], d["e2e"]["value"], "| 20 steps:", e["value"], e["e2e"]["value"], "| launches", d.get("gpu_launches"))
for k, w in d["workloads"].items():
    print(k, "value %.4g e2e %.4g" % (w["value"], w["e2e"]["value"]), {a: round(b, 4) for a, b in w["kernel_ms"].items()},
          "cpu %.4g" % w["cpu_baseline"]["value"], "fp64 frac %.4f" % w["fp64"]["frac"])
print("c5", d["c5"]["value"], d["c5"]["vecenv"]["value"], d["c5"]["autoreset_cost"]["overhead_frac"], d["c5"]["reset_all_ms"])
print("c1", d["c1"]["seconds_per_step"], d["c1"]["cpu_baseline"]["seconds_per_step"])
print("plan", {k: (v["iterations_per_s"], v["cpu_baseline"]["iterations_per_s"]) for k, v in d["plan"].items()})
print("fd", d["fd"]["ms_per_call"], d["fd"]["cpu_baseline"]["ms_per_call"])
print("ref", r["value"], {k: v["value"] for k, v in r["workloads"].items()})
print("roofline", d["roofline"])
print("clocks", d.get("clocks"))
for f in ("bench.json", "bench20.json", "bench_ref.json", "pytest_gpu.log", "smoke.log", "C3_launches.csv", "settled_launches.csv"):
    shutil.copy(f"{go}/{tag}_{f}", pr)
for name in ("C3", "settled"):
    out = f"{pr}/{tag}_{name}_ncu_full_summary.csv"
    args = [sys.executable, f"{root}/tools/dev/ncu_summary.py", f"{go}/{tag}_{name}_step.ncu-rep", out] + ([f"{pr}/traffic.json"] if name == "C3" else [])
    subprocess.run(args, capture_output=True)
    rows = list(csv.reader(open(out)))
    ix = {h: i for i, h in enumerate(rows[0])}; seen = set(); o = []
    for rr in rows[2:]:
        k = rr[0].split("(")[0].replace("void ", "")
        if k in seen: continue
        seen.add(k)
        o.append("%s %.0f us (warps %.0f%% issue %.0f%% fp64 %.0f%% inst %.1fM)" % (
            k, float(rr[ix["gpu__time_duration.sum"]]), float(rr[ix["sm__warps_active.avg.pct_of_peak_sustained_active"]]),
            float(rr[ix["smsp__issue_active.avg.pct_of_peak_sustained_active"]]),
            float(rr[ix["sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"]]), float(rr[ix["smsp__inst_executed.sum"]]) / 1e6))
    print(name, " | ".join(o))
