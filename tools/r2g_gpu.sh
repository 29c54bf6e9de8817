mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-cpu --no-configs"
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
DCEGM_TRACE=1 $B > gpurun_out/r2g_one.json 2> gpurun_out/r2g_one.err; tail -19 gpurun_out/r2g_one.err
python -c "
import json; d=json.load(open('gpurun_out/r2g_one.json')); print('RESULT', d['value'], d['e2e']['value'], d['e2e']['full_xy']['value'])"
