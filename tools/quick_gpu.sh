#!/bin/bash
# quick GPU check used during kernel work: F parity tests + short C3 bench line (value only)
python -m pytest tests/test_gpu_parity_f.py -x -q 2>&1 | tail -2
python bench.py --no-cpu --steps 3 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('value',round(d['value']),'e2e',round(d['e2e']['value']),'ms',round(d['ms_per_step'],2))"
