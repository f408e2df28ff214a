#!/bin/bash
T=${1:-r2k}
python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.txt
tail -5 gpurun_out/${T}_pytest.txt
python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/${T}_bench_c3.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_c3.json')); print(d['value'], d['ms_per_step'], d['stage_ms'], d['e2e']['value'], d['e2e_topk']['value'])"
for c in C2 C5; do python bench.py --config $c --steps 5 --warmup 3 --no-cpu --no-extras --no-e2e > gpurun_out/${T}_bench_$c.json 2>> gpurun_out/${T}_bench.err; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_$c.json')); print('$c', d['value'], d['ms_per_step'], d['stage_ms'])"; done
