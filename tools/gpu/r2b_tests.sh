#!/bin/bash
# full GPU test suite, then the op-count baseline
T=${1:-r2b}
python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.txt
tail -5 gpurun_out/${T}_pytest.txt
bash tools/gpu/r2a_baseline.sh $T
