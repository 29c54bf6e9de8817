# round-1 capture G (final build of the round): full-set ncu captures of the dominant kernels, each after the same
# command exited 0 without ncu, plus launch lists
set -x
python bench.py --steps 2 --warmup 3 --no-cpu --batch 1184 > gpurun_out/r1g_pre.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:solve_f_kernel -s 3 -c 1 -o gpurun_out/r1g_solve_f python bench.py --steps 2 --warmup 3 --no-cpu --batch 1184 > gpurun_out/r1g_ncu1.log 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r1g_pre2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1g_launches_bench_c3.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r1g_ncu2.log 2>&1
M=sm__inst_executed_pipe_fp64.sum,smsp__inst_executed.sum,gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
ncu --metrics $M --clock-control none -k regex:solve_f_kernel -s 3 -c 1 --csv --log-file gpurun_out/r1g_solve_f_metrics.csv python bench.py --steps 2 --warmup 3 --no-cpu --batch 1184 > gpurun_out/r1g_ncu2b.log 2>&1
python tools/sim_run.py 592 20000 g > gpurun_out/r1g_pre3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:sim_kernel -s 1 -c 1 -o gpurun_out/r1g_sim_g python tools/sim_run.py 592 20000 g > gpurun_out/r1g_ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:solve_g_kernel -s 1 -c 1 -o gpurun_out/r1g_solve_g python tools/sim_run.py 592 2000 g > gpurun_out/r1g_ncu3b.log 2>&1
python tools/config_runs.py c4 148 > gpurun_out/r1g_pre4.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:big_egm_kernel -s 20 -c 1 -o gpurun_out/r1g_big_egm python tools/config_runs.py c4 148 > gpurun_out/r1g_ncu4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:big_env_kernel -s 20 -c 1 -o gpurun_out/r1g_big_env python tools/config_runs.py c4 148 > gpurun_out/r1g_ncu5.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1g_launches_c4.csv python tools/config_runs.py c4 148 > gpurun_out/r1g_ncu6.log 2>&1
cat gpurun_out/r1g_pre4.log
