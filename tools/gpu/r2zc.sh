#!/bin/bash
python -m pytest tests/test_api_sweep.py tests/test_random_sweep.py -m gpu -q > gpurun_out/r2zc_pytest.txt 2>&1; tail -c 2500 gpurun_out/r2zc_pytest.txt
python bench.py --config C5 --precision fp32 --steps 5 --warmup 3 --no-cpu --no-extras > gpurun_out/r2zc_bench_C5_fp32.json 2>>gpurun_out/r2zc.err
python -c "
import json; d=json.loads(open('gpurun_out/r2zc_bench_C5_fp32.json').read().strip().splitlines()[-1]); print('C5 fp32', d['ms_per_step'], d['stage_ms'])"
