# A/B of the simulator: usage ab_sim.sh <tag> [lib ...]; first the sim tests on the default build
mkdir -p gpurun_out
tag=$1; shift
python -m pytest tests/test_gpu_sim.py tests/test_gpu_parity_g_big.py tests/test_gpu_sweep.py -x -q 2>&1 | tail -3
run() { for k in g f; do python tools/sim_run.py 2368 100000 $k | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('RESULT', '$1', d['kind'], round(d['sim_s']*1e3,1), 'ms', round(d['agent_periods_per_s']/1e9,2), 'G/s')" | tee -a gpurun_out/ab_sim_$tag.txt; done; }
run base
for lib in "$@"; do
  export DCEGM_CUDA_LIB=$PWD/dcegm.jl_b200/csrc/$lib
  run "$lib"
  unset DCEGM_CUDA_LIB
done
