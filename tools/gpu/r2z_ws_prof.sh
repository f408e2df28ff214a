#!/bin/bash
# per-line profile of the warp-specialised toppra kernel on C2 (its default there) and, forced, on C5
export SCB_TOPPRA_WS=1
for c in C2 C5; do
CMD="python bench.py --config $c --steps 2 --warmup 3 --no-e2e --no-cpu --no-extras"
$CMD > gpurun_out/r2z_plain_$c.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"toppra_ws_kernel" -s 2 -c 1 -o gpurun_out/r2z_ws_$c $CMD > gpurun_out/r2z_ncu_$c.log 2>&1
tail -2 gpurun_out/r2z_ncu_$c.log
done
