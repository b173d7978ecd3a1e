"""Top stall-sample SASS lines (and stall-reason columns) from `ncu --page source --csv` of an .ncu-rep."""
import csv, subprocess, sys
path = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]
si = hdr.index("# Samples"); src = hdr.index("Source"); ie = hdr.index("Instructions Executed")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") or h.lower().startswith("stall")]
data = []
for r in rows[hi + 1:]:
    try:
        data.append((int(r[si]), r))
    except Exception:
        pass
tot = sum(d[0] for d in data) or 1
print("total samples", tot, "instructions", len(data))
if stall_cols:
    agg = {hdr[i]: sum(int(r[i] or 0) for _, r in data if r[i].isdigit()) for i in stall_cols}
    print({k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:12]})
for idx, (n, r) in enumerate(data):
    pass
order = sorted(range(len(data)), key=lambda i: -data[i][0])[:top]
for i in sorted(order):
    n, r = data[i]
    st = {hdr[c][6:]: int(r[c]) for c in stall_cols if r[c].isdigit() and int(r[c]) > 0}
    st = dict(sorted(st.items(), key=lambda kv: -kv[1])[:3])
    print(f"{i:5d} {n:7d} {100*n/tot:5.1f}%  exec {r[ie]:>8s}  {r[src].strip()[:80]:80s} {st}")
