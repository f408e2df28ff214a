#!/bin/bash
python -m pytest tests -m gpu -q -x > gpurun_out/r2za_pytest.txt 2>&1; tail -3 gpurun_out/r2za_pytest.txt
for c in C2 C5; do
  python bench.py --config $c --steps 5 --warmup 3 --no-cpu --no-extras > gpurun_out/r2za_bench_$c.json 2>>gpurun_out/r2za.err
  python -c "
import json; d=json.loads(open('gpurun_out/r2za_bench_$c.json').read().strip().splitlines()[-1]); print('$c', d['ms_per_step'], d['stage_ms'], d['roofline']['path']['fp_frac'])"
done
python bench.py --config C5 --precision fp32 --steps 5 --warmup 3 --no-cpu --no-extras > gpurun_out/r2za_bench_C5_fp32.json 2>>gpurun_out/r2za.err
python bench.py --steps 10 --warmup 5 --no-cpu --no-extras > gpurun_out/r2za_bench_c3.json 2>>gpurun_out/r2za.err
python -c "
import json; d=json.loads(open('gpurun_out/r2za_bench_c3.json').read().strip().splitlines()[-1]); print('C3', d['ms_per_step'], d['stage_ms'])"
