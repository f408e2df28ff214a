#!/bin/bash
# baseline of the round-1 kernels: executed FP64 op counts per kernel (ncu), fresh C2 / C5 records
M=smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
T=${1:-r2a}
python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/${T}_plain.log 2>&1 && \
ncu --metrics $M --clock-control none -s 12 -c 10 --csv --log-file gpurun_out/${T}_ops.csv python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/${T}_ncu.log 2>&1
for c in C2 C5; do python bench.py --config $c --steps 5 --warmup 3 --no-cpu > gpurun_out/${T}_$c.json 2> gpurun_out/${T}_$c.err; done
python bench.py --config C5 --precision fp32 --steps 5 --warmup 3 --no-cpu > gpurun_out/${T}_C5_fp32.json 2> gpurun_out/${T}_C5_fp32.err
nproc; free -g | head -2
