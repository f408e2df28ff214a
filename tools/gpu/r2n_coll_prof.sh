#!/bin/bash
CMD="python bench.py --obstacles 24 --steps 2 --warmup 3 --no-e2e --no-cpu --no-extras"
$CMD > gpurun_out/r2n_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"path_lane_kernel" -s 2 -c 1 -o gpurun_out/r2n_coll $CMD > gpurun_out/r2n_ncu.log 2>&1
tail -3 gpurun_out/r2n_ncu.log
