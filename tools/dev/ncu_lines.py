"""Per-source-line samples / instructions from `ncu --page source --print-source cuda,sass --csv` output."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 50
fn = None; out = []
for r in rows:
    if r and r[0] == "File Path": fn = r[1].split("/")[-1]; continue
    if len(r) < 10 or r[0] in ("Line No", ""): continue
    try: out.append((int(r[6]), int(r[7]), fn, int(r[0]), r[1].strip()[:120]))
    except ValueError: pass
ts, ti = sum(o[0] for o in out), sum(o[1] for o in out)
print("total samples", ts, "total warp-inst", ti)
out.sort(reverse=True)
acc = 0
for o in out[:top]:
    acc += o[0]
    print(f"{100*o[0]/ts:5.1f}% {100*acc/ts:5.1f}% inst {100*o[1]/ti:5.1f}%  {o[2]}:{o[3]}  {o[4]}")
