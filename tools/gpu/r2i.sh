#!/bin/bash
T=${1:-r2i}
python -m pytest tests/test_scale_parity.py tests/test_collision.py tests/test_toppra_ppoly.py -m gpu -q > gpurun_out/${T}_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.txt
tail -8 gpurun_out/${T}_pytest.txt
python bench.py --obstacles 24 --steps 6 --warmup 3 --no-cpu --no-extras > gpurun_out/${T}_bench_c3_obstacles.json 2> gpurun_out/${T}_bench.err; echo "obst rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_c3_obstacles.json')); print(d['ms_per_step'], d['stage_ms'], d['run']['statuses'])"
