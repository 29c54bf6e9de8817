for i in 1 2; do
for lib in libdcegm_cuda.so libdcegm_cuda_varA.so libdcegm_cuda_varB.so; do
DCEGM_CUDA_LIB=$PWD/dcegm.jl_b200/csrc/$lib python bench.py --no-cpu --steps 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$lib','value',round(d['value']),'ms',round(d['ms_per_step'],2))"
done; done
for lib in libdcegm_cuda.so libdcegm_cuda_varA.so libdcegm_cuda_varB.so; do DCEGM_CUDA_LIB=$PWD/dcegm.jl_b200/csrc/$lib python tools/config_runs.py c5 2000 | cut -c1-160; done
