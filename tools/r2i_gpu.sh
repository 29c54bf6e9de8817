mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity_f_big.py tests/test_gpu_compact.py tests/test_gpu_sim.py -x -q 2>&1 | tail -3
python tools/config_runs.py c4 256 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('RESULT c4', round(d['kernel_ms'],1), d['launches'], d['solved'])"
DCEGM_BIG_PARTS=8 python tools/config_runs.py c4 256 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('RESULT c4 parts8', round(d['kernel_ms'],1), d['launches'], d['solved'])"
DCEGM_BIG_PARTS=2 python tools/config_runs.py c4 256 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('RESULT c4 parts2', round(d['kernel_ms'],1), d['launches'], d['solved'])"
python tools/c4_e2e.py 256 | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('RESULT', d['config'][:40], round(d['e2e_ms'],1))"
