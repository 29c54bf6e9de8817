set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2c_pytest.log 2>&1; tail -5 gpurun_out/r2c_pytest.log
DCEGM_TRACE=1 python bench.py --steps 3 --warmup 3 --no-cpu --no-configs > gpurun_out/r2c_trace.json 2> gpurun_out/r2c_trace.err; tail -14 gpurun_out/r2c_trace.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2c_trace.json'))
print(d['value'], d['e2e']['value'], d['e2e']['full_xy']['value'], d['e2e']['d2h_gbs'])
PY
