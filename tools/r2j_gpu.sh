mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2j_pytest.log 2>&1; tail -3 gpurun_out/r2j_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/r2j_bench.json 2> gpurun_out/r2j_bench.err; tail -c 400 gpurun_out/r2j_bench.json; echo
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2j_bench_ref.json 2>/dev/null; head -c 300 gpurun_out/r2j_bench_ref.json; echo
bash tools/ncu_round2.sh r2j > gpurun_out/r2j_ncu_script.log 2>&1; tail -3 gpurun_out/r2j_ncu_script.log
