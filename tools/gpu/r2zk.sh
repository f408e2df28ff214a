#!/bin/bash
python -m pytest tests/test_gpu_parity.py -m gpu -q -k "pair_mapping" 2>&1 | tail -12
run() { python bench.py --config ${CFG:-C3} --steps 10 --warmup 5 --no-cpu --no-extras --no-e2e 2>>gpurun_out/r2zk.err | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['ms_per_step'],3), round(d['value']), {k: round(v,3) for k,v in d['stage_ms'].items()})"; }
SCB_PATH_PAIR=0 run "C3 pair off"
run "C3 pair default"
CFG=C2 SCB_PATH_PAIR=0 run "C2 pair off"
CFG=C2 run "C2 pair default"
