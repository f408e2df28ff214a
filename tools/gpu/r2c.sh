#!/bin/bash
# GPU test suite (no -x: see everything that fails), then the bench line of both arms
T=${1:-r2c}
python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.txt
tail -15 gpurun_out/${T}_pytest.txt
python bench.py --steps 10 --warmup 3 > gpurun_out/${T}_bench_c3.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${T}_bench_ref.json 2>> gpurun_out/${T}_bench.err; echo "ref rc=$?"
tail -c 600 gpurun_out/${T}_bench.err
