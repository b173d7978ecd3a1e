# same-box A/B of library variants: tools/ab_bench.sh "<name> <name> ..."   (build/variants/lib_<name>.so, two rounds each)
for r in 1 2; do
for v in $1; do
  cp build/variants/lib_$v.so eventgpt_b200/lib/libegpt_b200.so
  echo "== $v"
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
j = json.loads(sys.stdin.read().strip().splitlines()[-1]); k = j['kernel_ms_per_step']
print('value', round(j['value'],1), 'ms', round(j['ms_per_step'],2), 'clk', j['clocks']['sm_mhz'], 'gemm', round(k['gemm'],2), 'attn', round(k['attention'],2), 'elem', round(k.get('elementwise',0),2), 'TF', round(j['roofline']['achieved']))"
done
done
