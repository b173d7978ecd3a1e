# The GPU pass behind the round's final numbers: bash tools/final_pass.sh (under gpurun; writes into gpurun_out/)
set -x
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_final3.json 2> gpurun_out/r2_bench_final3.err; head -c 240 gpurun_out/r2_bench_final3.json; echo
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_bench_ref3.json 2>/dev/null; head -c 200 gpurun_out/r2_bench_ref3.json; echo
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
