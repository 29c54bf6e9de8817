# round-2 ncu captures of the final build (every ncu command after the same command exited 0 without ncu):
# launch list of the default bench, counters + full-set capture of solve_f_kernel, the compact encoder kernels, C4 launch list
set -x
mkdir -p gpurun_out
T=${1:-r2f}
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-configs"
$B > gpurun_out/${T}_pre.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/${T}_launches_bench_c3.csv $B > gpurun_out/${T}_ncu1.log 2>&1
M=sm__inst_executed_pipe_fp64.sum,smsp__inst_executed.sum,gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
$B --batch 1184 > gpurun_out/${T}_pre2.log 2>&1 && \
ncu --metrics $M --clock-control none -k regex:solve_f_kernel -s 3 -c 1 --csv --log-file gpurun_out/${T}_solve_f_metrics.csv $B --batch 1184 > gpurun_out/${T}_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:solve_f_kernel -s 3 -c 1 -o gpurun_out/${T}_solve_f $B --batch 1184 > gpurun_out/${T}_ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:compact_count_kernel -s 2 -c 1 -o gpurun_out/${T}_compact_count $B --batch 1184 > gpurun_out/${T}_ncu4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:compact_write_kernel -s 2 -c 1 -o gpurun_out/${T}_compact_write $B --batch 1184 > gpurun_out/${T}_ncu5.log 2>&1
python tools/config_runs.py c4 148 > gpurun_out/${T}_pre6.log 2>&1 && \
DCEGM_BIG_PARTS=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/${T}_launches_c4.csv python tools/config_runs.py c4 148 > gpurun_out/${T}_ncu6.log 2>&1
ls -la gpurun_out/${T}_*
