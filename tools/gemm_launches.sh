# per-launch GEMM durations of two tower layers inside bench.py for library variants: tools/gemm_launches.sh "<name> ..."
for v in $1; do
  cp build/variants/lib_$v.so eventgpt_b200/lib/libegpt_b200.so
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:gemm2 -s 21 -c 8 --csv --log-file gpurun_out/gl_$v.csv \
      python bench.py --steps 1 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
  python - "$v" <<'PY'
import csv, sys
v = sys.argv[1]
rows = [r for r in csv.reader(open(f"gpurun_out/gl_{v}.csv")) if len(r) > 5]
h = rows[0]; ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
print(v, " ".join(f"{r[ki].split('<')[1].split('>')[0]}:{float(r[vi].replace(',',''))/ (1e3 if r[ui]=='ns' else 1):.0f}" for r in rows[1:]))
PY
done
