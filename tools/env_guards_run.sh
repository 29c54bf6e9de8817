# Shared-memory guard check of the envelope code (stand-in for compute-sanitizer memcheck, closed on this pool):
# the library built with -DDC_ENV_GUARDS puts guard bytes behind every array of the envelope scratch and reports a
# written guard as status bit 15; the parity tests (status compared with the oracle's) then run on that build.
#   make -C dcegm.jl_b200/csrc variant VARIANT_FLAGS="-DDC_ENV_GUARDS" && mv .../libdcegm_cuda_var.so .../libdcegm_cuda_var_guards.so
#   (positive control: the same with -DDC_ENV_GUARD_SELFTEST as libdcegm_cuda_var_guardst.so)
mkdir -p gpurun_out
L=$PWD/dcegm.jl_b200/csrc
DCEGM_CUDA_LIB=$L/libdcegm_cuda_var_guards.so python -m pytest tests/test_gpu_parity_f.py tests/test_gpu_parity_f_big.py tests/test_gpu_parity_g_minimal.py tests/test_gpu_parity_g_big.py tests/test_gpu_compact.py tests/test_gpu_sim.py -m gpu -q 2>&1 | tail -3
DCEGM_CUDA_LIB=$L/libdcegm_cuda_var_guards.so python - <<'PY'
import numpy as np
from bench_configs import c3_params, c4_params, c5_params
from dcegm_jl_b200 import _abi
from dcegm_jl_b200.solver import Solver
with Solver(0) as s:
    for name, model, arr in (("C3 x4096", _abi.MODEL_F, c3_params(4096)), ("C5 x8192", _abi.MODEL_G, c5_params(8192, 1000)),
                             ("C4 x148", _abi.MODEL_F, c4_params(148))):
        s.solve(model, arr, fetch=False)
        st = s.status().astype(np.int64)
        print("guards build,", name, ": models with a written guard (status bit 15):", int(((st >> 15) & 1).sum()), "of", len(st))
PY
DCEGM_CUDA_LIB=$L/libdcegm_cuda_var_guardst.so python - <<'PY'
import numpy as np
from dcegm_jl_b200 import Param, _abi
from dcegm_jl_b200.solver import Solver
with Solver(0) as s:
    for model, ps in ((_abi.MODEL_F, [Param(par=dict(na=200))] * 3), (_abi.MODEL_G, [Param(par=dict(na=100))] * 3)):
        st = s.solve(model, ps).status.astype(np.int64)
        print("positive control, model", model, "status bit 15 set in all models:", bool(((st >> 15) & 1).all()))
    st = s.solve(_abi.MODEL_F, [Param(par=dict(na=200))] * 2, fine_grid=True).status.astype(np.int64)
    print("positive control, fine-grid solver:", bool(((st >> 15) & 1).all()))
PY
