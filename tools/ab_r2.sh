# A/B of solve_f_kernel variants on the C3 bench (value only): usage ab_r2.sh <tag> [lib ...]
set -x
mkdir -p gpurun_out
tag=$1; shift
for lpt in 0 1; do
  DCEGM_LPT=$lpt python bench.py --steps 5 --warmup 3 --no-cpu --no-configs | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('RESULT base lpt=$lpt', round(d['value']), round(d['ms_per_step'],3), round(d['e2e']['value']))" | tee -a gpurun_out/ab_$tag.txt
done
for lib in "$@"; do
  DCEGM_CUDA_LIB=$PWD/dcegm.jl_b200/csrc/$lib python bench.py --steps 5 --warmup 3 --no-cpu --no-configs | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('RESULT $lib', round(d['value']), round(d['ms_per_step'],3), round(d['e2e']['value']))" | tee -a gpurun_out/ab_$tag.txt
done
