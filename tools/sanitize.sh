#!/bin/bash
# compute-sanitizer over the integer kernels and a 2-layer token call: memcheck, racecheck, synccheck.
# Usage (GPU box): bash tools/sanitize.sh > gpurun_out/sanitizer.log 2>&1
set -u
CS=$(command -v compute-sanitizer || echo /usr/local/cuda/bin/compute-sanitizer)
$CS --version | head -2
run() {   # tool, label, command...
  tool=$1; label=$2; shift 2
  echo "=== compute-sanitizer --tool $tool : $label"
  timeout 900 $CS --tool "$tool" --print-limit 5 "$@" > /tmp/cs.log 2>&1
  echo "exit code $?"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|Error:|hazard|passed|failed|error" /tmp/cs.log | tail -6
}
RASTER='tests/test_gpu_events.py -x -q -m gpu -k "bit_exact and (sample1 or adv_zipf or adv_n5 or synthetic_000003) or batched_ragged or generic_dtypes or device_clip or split"'
for tool in memcheck racecheck synccheck; do
  run $tool "raster / preprocess / split tests" bash -c "python -m pytest $RASTER"
done
run memcheck "no-allocation raster" python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu
run memcheck "gemm + attention unit tests, 2-layer fused tokens" bash -c 'python -m pytest tests/test_gpu_tower.py -x -q -m gpu -k "gemm_vs or attention_vs or fused_events or splice_bit"'
run racecheck "attention unit test, 2-layer fused tokens" bash -c 'python -m pytest tests/test_gpu_tower.py -x -q -m gpu -k "attention_vs or fused_events"'
run synccheck "2-layer fused tokens" bash -c 'python -m pytest tests/test_gpu_tower.py -x -q -m gpu -k "fused_events"'
