#!/bin/bash
# Times the attention kernel for every variant library under build/variants/ (developer aid for parameter sweeps).
for f in build/variants/lib_*.so; do
  cp "$f" eventgpt_b200/lib/libegpt_b200.so
  echo "== $f"; timeout 120 python tools/bringup.py attn 2>&1 | grep "F=160"
done
