#!/bin/bash
# Does capping the residency of the FP64-bound kernels let the other batch's latency-bound kernels run beside them?
run() { python bench.py --in-flight ${NFL:-2} --steps 12 --warmup 6 --no-cpu --no-extras --no-e2e 2>>gpurun_out/r2zf.err | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['ms_per_step'],3), round(d['value']), {k: round(v,3) for k,v in d['stage_ms'].items()})"; }
run "default"
SCB_PATH_SMEM_KB=60 run "path 3/SM"
SCB_PATH_SMEM_KB=60 SCB_ARC_SMEM_KB=80 run "path 3/SM arc 2/SM"
SCB_ARC_SMEM_KB=80 run "arc 2/SM"
SCB_PATH_SMEM_KB=80 run "path 2/SM"
NFL=3 SCB_PATH_SMEM_KB=60 run "path 3/SM, 3 in flight"
NFL=3 SCB_PATH_SMEM_KB=60 SCB_ARC_SMEM_KB=80 run "path 3 arc 2, 3 in flight"
NFL=4 SCB_PATH_SMEM_KB=60 SCB_ARC_SMEM_KB=80 run "path 3 arc 2, 4 in flight"
