#!/bin/bash
# GPU test suite, bench of both arms, the obstacle workload (collision term inside the optimiser), op counts of the final kernels
T=${1:-r2e}
python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.txt
tail -12 gpurun_out/${T}_pytest.txt
python bench.py --steps 10 --warmup 3 > gpurun_out/${T}_bench_c3.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${T}_bench_ref.json 2>> gpurun_out/${T}_bench.err; echo "ref rc=$?"
python bench.py --obstacles 24 --steps 6 --warmup 3 --no-cpu --no-extras > gpurun_out/${T}_bench_c3_obstacles.json 2>> gpurun_out/${T}_bench.err; echo "obst rc=$?"
tail -c 800 gpurun_out/${T}_bench.err
