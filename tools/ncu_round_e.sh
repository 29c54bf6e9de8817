set -x
python tools/dfma_rate.py > gpurun_out/r1e_dfma_rate.json 2>&1
M=sm__inst_executed_pipe_fp64.sum,smsp__inst_executed.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__warps_active.avg.per_cycle_active,launch__registers_per_thread
python bench.py --steps 2 --warmup 3 --no-cpu --batch 1184 > gpurun_out/r1e_pre.log 2>&1 && \
ncu --metrics $M --clock-control none -k regex:solve_f_kernel -s 3 -c 1 --csv --log-file gpurun_out/r1e_solve_f_metrics.csv python bench.py --steps 2 --warmup 3 --no-cpu --batch 1184 > gpurun_out/r1e_ncu1.log 2>&1
python tools/config_runs.py c4 64 > gpurun_out/r1e_c4_64.log 2>&1 && \
ncu --metrics $M --clock-control none -k regex:big_ -s 40 -c 8 --csv --log-file gpurun_out/r1e_big_metrics.csv python tools/config_runs.py c4 64 > gpurun_out/r1e_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:big_egm_kernel -s 20 -c 1 -o gpurun_out/r1e_big_egm python tools/config_runs.py c4 64 > gpurun_out/r1e_ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:big_env_kernel -s 20 -c 1 -o gpurun_out/r1e_big_env python tools/config_runs.py c4 64 > gpurun_out/r1e_ncu4.log 2>&1
cat gpurun_out/r1e_dfma_rate.json gpurun_out/r1e_c4_64.log
