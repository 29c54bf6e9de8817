# round-2 call B: compact delivery tests + the new bench line
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1; tail -5 gpurun_out/r2b_pytest.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; tail -c 3000 gpurun_out/r2b_bench.json; tail -5 gpurun_out/r2b_bench.err
DCEGM_TRACE=1 python bench.py --steps 2 --warmup 3 --no-cpu --no-configs > gpurun_out/r2b_trace.json 2> gpurun_out/r2b_trace.err; tail -60 gpurun_out/r2b_trace.err
