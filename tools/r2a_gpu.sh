# round-2 call A: GPU tests on the re-entry build, flag census of C3/C4/C5 on the GPU, compute-sanitizer logs
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; tail -3 gpurun_out/r2a_pytest.log
python bench.py --steps 3 --warmup 3 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; tail -c 600 gpurun_out/r2a_bench.json
python tools/flag_census.py c3 --backend gpu > gpurun_out/r2a_census_c3.json 2> gpurun_out/r2a_census_c3.examples
python tools/flag_census.py c5 --backend gpu > gpurun_out/r2a_census_c5.json 2> gpurun_out/r2a_census_c5.examples
python tools/flag_census.py c4 --backend gpu --n 256 --chunk 4 --examples 40 > gpurun_out/r2a_census_c4.json 2> gpurun_out/r2a_census_c4.examples
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool python tools/sanitize_run.py > gpurun_out/r2a_sanitizer_$tool.log 2>&1
  tail -4 gpurun_out/r2a_sanitizer_$tool.log
done
cat gpurun_out/r2a_census_c3.json gpurun_out/r2a_census_c5.json gpurun_out/r2a_census_c4.json
