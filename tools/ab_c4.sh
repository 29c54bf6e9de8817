# A/B of the fine-grid solver on C4 (256 models): usage ab_c4.sh <tag> [lib ...]
mkdir -p gpurun_out
tag=$1; shift
run() { python tools/config_runs.py c4 256 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('RESULT', '$1', round(d['kernel_ms'],1), d['launches'])" | tee -a gpurun_out/ab_c4_$tag.txt; }
run base
DCEGM_BIG_PARTS=8 run base_parts8
for lib in "$@"; do
  export DCEGM_CUDA_LIB=$PWD/dcegm.jl_b200/csrc/$lib
  DCEGM_BIG_PARTS=4 run "$lib parts4"
  DCEGM_BIG_PARTS=8 run "$lib parts8"
  DCEGM_BIG_PARTS=2 run "$lib parts2"
  unset DCEGM_CUDA_LIB
done
