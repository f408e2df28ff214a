#!/bin/bash
run() { python bench.py --config ${CFG:-C3} --in-flight ${NFL:-2} --steps 12 --warmup 6 --no-cpu --no-extras --no-e2e 2>>gpurun_out/r2zg.err | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['ms_per_step'],3), round(d['value']), {k: round(v,3) for k,v in d['stage_ms'].items()})"; }
SCB_PRIO=0 run "prio off"
run "prio on"
NFL=3 run "prio on, 3 in flight"
NFL=4 run "prio on, 4 in flight"
SCB_PATH_SMEM_KB=60 run "prio on, path 3/SM"
NFL=3 SCB_PATH_SMEM_KB=60 run "prio on, path 3/SM, 3 in flight"
NFL=3 SCB_PATH_SMEM_KB=60 SCB_ARC_SMEM_KB=80 run "prio on, path 3 arc 2, 3 in flight"
CFG=C5 NFL=4 SCB_PRIO=0 run "C5 prio off"
CFG=C5 NFL=4 run "C5 prio on"
CFG=C2 NFL=4 SCB_PRIO=0 run "C2 prio off"
CFG=C2 NFL=4 run "C2 prio on"
python -m pytest tests/test_gpu_parity.py tests/test_records_fetch.py -m gpu -q -x 2>&1 | tail -2
