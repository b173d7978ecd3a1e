#!/bin/bash
# per-window step time as a function of the batch (tuning aid: would running the tower on sub-batches keep activations in L2?)
for b in 4 8 16 32; do
  EGPT_BENCH_BATCH=$b timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
j = json.loads(sys.stdin.read().strip().splitlines()[-1])
k = j['kernel_ms_per_step']; b = j['config']['windows_per_gpu']
print('batch', b, 'windows/s', round(j['value'],1), 'ms/window', round(j['ms_per_step']/b,3), 'clk', j['clocks']['sm_mhz'], {n: round(v/b,3) for n,v in k.items()})"
done
