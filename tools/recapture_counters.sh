# re-captures the ncu counters of solve_f_kernel after a change of its sources: usage recapture_counters.sh <tag>
mkdir -p gpurun_out
T=${1:-r2k}
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-configs"
python -m pytest tests/test_gpu_parity_f.py tests/test_gpu_parity_f_big.py tests/test_gpu_compact.py -x -q 2>&1 | tail -2
M=sm__inst_executed_pipe_fp64.sum,smsp__inst_executed.sum,gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
$B --batch 1184 > gpurun_out/${T}_pre2.log 2>&1 && \
ncu --metrics $M --clock-control none -k regex:solve_f_kernel -s 3 -c 1 --csv --log-file gpurun_out/${T}_solve_f_metrics.csv $B --batch 1184 > gpurun_out/${T}_ncu2.log 2>&1
python bench.py --no-cpu --no-configs --steps 10 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('RESULT', round(d['value']), round(d['e2e']['value']))"
