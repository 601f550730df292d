"""Per-source-line hot spots of one kernel in an .ncu-rep: python tools/ncu_lines.py file.ncu-rep [kernel-regex] [top]
Prints line, warp instructions executed, average active threads, stall samples and the source text."""
import csv, io, subprocess, sys
rep = sys.argv[1]; kre = sys.argv[2] if len(sys.argv) > 2 else ""; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
cmd = ["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"]
if kre:
    cmd += ["-k", "regex:" + kre]
out = subprocess.run(cmd, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = None; acc = {}
for r in rows:
    if len(r) > 5 and r[0] == "Line No":
        hdr = r; continue
    if hdr is None or len(r) != len(hdr) or not r[0].isdigit():
        continue
    d = dict(zip(hdr, r))
    try:
        ie = int(d["Instructions Executed"]); te = int(d["Thread Instructions Executed"]); ss = int(d["# Samples"])
    except (ValueError, KeyError):
        continue
    a = acc.setdefault(int(r[0]), [0, 0, 0, r[1]])
    a[0] += ie; a[1] += te; a[2] += ss
tot_i = sum(a[0] for a in acc.values()); tot_s = sum(a[2] for a in acc.values())
print(f"total warp inst {tot_i}, samples {tot_s}")
for ln, a in sorted(acc.items(), key=lambda kv: -kv[1][2])[:top]:
    print(f"{ln:5d} inst {100*a[0]/max(tot_i,1):5.1f}%  thr/inst {a[1]/max(a[0],1):5.1f}  samples {100*a[2]/max(tot_s,1):5.1f}%  {a[3].strip()[:110]}")
