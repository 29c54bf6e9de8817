set -x
mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-cpu --no-configs"
python -m pytest tests/test_gpu_compact.py tests/test_gpu_parity_f.py tests/test_gpu_parity_f_big.py -x -q 2>&1 | tail -3
DCEGM_TRACE=1 $B > gpurun_out/r2g_one.json 2> gpurun_out/r2g_one.err; tail -13 gpurun_out/r2g_one.err
python -c "
import json; d=json.load(open('gpurun_out/r2g_one.json')); print('RESULT', d['value'], d['e2e']['value'], d['e2e']['d2h_bytes_per_step'], d['e2e']['expand_all_on_host'])"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:compact -c 12 --csv --log-file gpurun_out/r2g_compact_launches.csv $B > /dev/null 2>&1
grep compact gpurun_out/r2g_compact_launches.csv | awk -F'","' '{print substr($5,1,22), $9, $NF}' | head -12
