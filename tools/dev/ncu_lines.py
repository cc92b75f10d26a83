"""Per-source-line stall samples from `ncu -i rep --page source --csv --print-source cuda,sass [--kernel-name ...]`.
usage: ncu_lines.py file.csv <function substring> [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
fn = func = None
hdr = None
out = []
for r in rows:
    if not r: continue
    if r[0] == "File Path": fn = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": func = r[1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or want not in (func or "") or len(r) < len(hdr) or r[2] != "-": continue
    try:
        d = dict(zip(hdr[4:], r[4:]))
        out.append((int(d["# Samples"]), int(d["Instructions Executed"]), int(d["stall_no_inst"]), int(d["stall_long_sb"]),
                    int(d["stall_short_sb"]), int(d["stall_wait"]), fn, int(r[0]), r[1].strip()[:90]))
    except (ValueError, KeyError):
        pass
ts, ti = sum(o[0] for o in out), sum(o[1] for o in out)
print("total samples", ts, "warp-inst", ti, "no_inst", sum(o[2] for o in out), "long_sb", sum(o[3] for o in out),
      "short_sb", sum(o[4] for o in out), "wait", sum(o[5] for o in out))
out.sort(reverse=True)
acc = 0
for o in out[:top]:
    acc += o[0]
    print(f"{100*o[0]/ts:5.1f}% {100*acc/ts:5.1f}% inst {100*o[1]/max(ti,1):5.1f}% noinst {o[2]:6d} lsb {o[3]:5d} ssb {o[4]:5d} wait {o[5]:5d}  {o[6]}:{o[7]}  {o[8]}")
