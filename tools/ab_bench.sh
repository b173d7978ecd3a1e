for r in 1 2; do
for v in old new; do
  cp build/variants/lib_$v.so eventgpt_b200/lib/libegpt_b200.so
  echo "== $v"
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
j = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('value', round(j['value'],1), 'ms', round(j['ms_per_step'],2), 'clk', j['clocks']['sm_mhz'], 'gemm', round(j['kernel_ms_per_step']['gemm'],2), 'attn', round(j['kernel_ms_per_step']['attention'],2), 'elem', round(j['kernel_ms_per_step']['elementwise'],2), 'TF', round(j['roofline']['achieved']))"
done
done
