# The GPU pass behind the round's final numbers: bash tools/final_pass.sh (under gpurun; writes into gpurun_out/)
set -x
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_final4.json 2> gpurun_out/r2_bench_final4.err; head -c 240 gpurun_out/r2_bench_final4.json; echo
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_bench_ref4.json 2>/dev/null; head -c 200 gpurun_out/r2_bench_ref4.json; echo
python bench.py --ingest --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_ingest4.json 2>/dev/null
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 1300 --csv --log-file gpurun_out/r2_launches4.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2_ncu_ll4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"gemm2_bf16|attention_tc" -s 16 -c 10 -o gpurun_out/r2_layer_final4 -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_lf4.log 2>&1
timeout 120 python tools/run_attn.py
timeout 120 python tools/run_raster.py
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
