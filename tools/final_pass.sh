set -x
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_final2.json 2> gpurun_out/r2_bench_final2.err; tail -c 600 gpurun_out/r2_bench_final2.json | head -c 300
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_bench_ref2.json 2>/dev/null
python bench.py --ingest --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_ingest2.json 2>/dev/null
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 1300 --csv --log-file gpurun_out/r2_launches2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2_ncu_ll2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"gemm2_bf16|attention_tc" -s 16 -c 10 -o gpurun_out/r2_layer_final2 -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_lf2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:resample_ -s 4 -c 2 -o gpurun_out/r2_pp2 -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_pp2.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
