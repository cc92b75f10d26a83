"""Per-source-line share of the EXECUTED warp-instructions of a kernel (and of its stall samples) from
`ncu -i rep --page source --csv --print-source cuda,sass --kernel-name ...`.
usage: ncu_inst.py file.csv <function substring> [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
fn = func = None; hdr = None; out = {}
for r in rows:
    if not r: continue
    if r[0] == "File Path": fn = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": func = r[1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or want not in (func or "") or len(r) < len(hdr) or r[2] != "-": continue
    try:
        d = dict(zip(hdr[4:], r[4:]))
        o = out.setdefault((fn, int(r[0])), [0, 0, r[1].strip()[:100]])
        o[0] += int(d["Instructions Executed"]); o[1] += int(d["# Samples"])
    except (ValueError, KeyError):
        pass
ti = sum(v[0] for v in out.values()); ts = sum(v[1] for v in out.values())
print("total warp-instructions", ti, "samples", ts)
acc = 0
for k, v in sorted(out.items(), key=lambda x: -x[1][0])[:top]:
    acc += v[0]
    print(f"{100*v[0]/ti:5.1f}% {100*acc/ti:5.1f}%  samples {100*v[1]/max(ts,1):5.1f}%  {k[0]}:{k[1]}  {v[2]}")
