"""Collects the bench.py JSON lines of this round (gpurun_out/r2_bench_*.json) into a markdown table + the raw lines.
Usage: python tools/bench_history.py > profiles/r02_bench_history.md"""
import glob, json, os, re
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rows = []
for p in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "r2_bench_*.json"))):
    try:
        line = [l for l in open(p).read().splitlines() if l.startswith("{")][-1]
        rows.append((os.path.basename(p), json.loads(line), line))
    except Exception:
        pass
print("# Round 2 — bench.py lines (one `gpurun` box each; boxes differ by a few percent under the 1 000 W power cap)\n")
print("| file | impl / config | GPUs | windows (or Mevents)/s | ms/step | e2e | SM MHz | GEMM TF/s (frac) | raster frac | kernel ms/step |")
print("|---|---|---:|---:|---:|---:|---:|---:|---:|---|")
for name, j, _ in rows:
    cfg = re.search(r"configs\[(\d)\]", j.get("config", {}).get("workload", ""))
    k = j.get("kernel_ms_per_step", {})
    r, rr = j.get("roofline", {}), j.get("raster_roofline", {})
    print(f"| `{name}` | {j.get('impl', 'ours')} configs[{cfg.group(1) if cfg else '?'}] | {j.get('n_gpus')} | {j['value']:.1f} | {j['ms_per_step']:.2f} | "
          f"{j.get('e2e', {}).get('value', 0):.1f} | {(j.get('clocks') or {}).get('sm_mhz', '')} | "
          f"{r.get('achieved', 0):.0f} ({r.get('frac', 0):.2f}) | {rr.get('frac', 0) or 0:.3f} | "
          + ", ".join(f"{a} {b:.2f}" for a, b in k.items()) + " |")
print("\n## Raw lines\n")
for name, _, line in rows:
    print(f"`{name}`\n\n```json\n{line}\n```\n")
