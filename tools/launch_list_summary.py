"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.
Usage: python tools/launch_list_summary.py gpurun_out/launches.csv "<state of the code>" > profiles/rNN_launch_list_summary.md
(also writes profiles/rNN_launches.csv.gz next to it when a third argument names the output prefix)."""
import collections, csv, gzip, shutil, sys

path = sys.argv[1]
state = sys.argv[2] if len(sys.argv) > 2 else ""
rows = [r for r in csv.reader(open(path)) if len(r) > 5]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[1:]:
    v = float(r[vi].replace(",", ""))
    us = {"ns": v / 1e3, "us": v, "usecond": v, "ms": v * 1e3, "msecond": v * 1e3}.get(r[ui], v / 1e3)
    name = r[ki].split("(")[0].replace("void ", "")
    tot[name] += us
    cnt[name] += 1
T = sum(tot.values())
print("# ncu launch list of `python bench.py --steps 2 --warmup 3 --no-cpu-baseline` (first 1300 launches)\n")
print("`ncu --metrics gpu__time_duration.sum --clock-control none -c 1300` after the same command exited 0 without ncu (same gpurun call).")
print(f"Serialised, cold-cache: compare SHARES.  State of the code: {state}\n")
print("| kernel | launches | total us | share |\n|---|---:|---:|---:|")
for k, v in tot.most_common():
    print(f"| `{k}` | {cnt[k]} | {v:.0f} | {100 * v / T:.1f}% |")
print(f"\ntotal {T / 1000:.1f} ms over {sum(cnt.values())} launches")
if len(sys.argv) > 3:
    with open(path, "rb") as f, gzip.open(sys.argv[3] + "_launches.csv.gz", "wb") as g:
        shutil.copyfileobj(f, g)
