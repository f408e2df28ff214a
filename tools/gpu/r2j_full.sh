#!/bin/bash
# Full record of the build on one B200: test suite, both bench arms, other configs, obstacle workload, ncu op counts,
# ncu launch list and one --set full capture of the step's kernels.
T=${1:-r2j}
python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.txt
tail -6 gpurun_out/${T}_pytest.txt
python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench_c3.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${T}_bench_ref.json 2>> gpurun_out/${T}_bench.err; echo "ref rc=$?"
python bench.py --obstacles 24 --steps 6 --warmup 3 --no-cpu --no-extras > gpurun_out/${T}_bench_c3_obstacles.json 2>> gpurun_out/${T}_bench.err; echo "obst rc=$?"
for c in C2 C5; do python bench.py --config $c --steps 5 --warmup 3 --no-cpu --no-extras > gpurun_out/${T}_bench_$c.json 2>> gpurun_out/${T}_bench.err; echo "$c rc=$?"; done
python bench.py --config C5 --precision fp32 --steps 5 --warmup 3 --no-cpu --no-extras > gpurun_out/${T}_bench_C5_fp32.json 2>> gpurun_out/${T}_bench.err
python bench.py --spline ref_cubic --steps 5 --warmup 3 --no-cpu --no-extras > gpurun_out/${T}_bench_c3_ref_cubic.json 2>> gpurun_out/${T}_bench.err
CMD="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --no-extras"
M=smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
$CMD > gpurun_out/${T}_plain.log 2>&1 && \
ncu --metrics $M --clock-control none -s 12 -c 10 --csv --log-file gpurun_out/${T}_ops.csv $CMD > gpurun_out/${T}_ncu_ops.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -s 12 -c 40 --csv --log-file gpurun_out/${T}_launches.csv $CMD > gpurun_out/${T}_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"path_pair_kernel|path_lane_kernel|arc_kernel|toppra_fast_kernel|param_kernel" -s 8 -c 4 -o gpurun_out/${T}_full $CMD > gpurun_out/${T}_ncu_full.log 2>&1
ls -la gpurun_out | tail -20
