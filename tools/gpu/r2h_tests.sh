#!/bin/bash
T=${1:-r2h}
python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.txt
tail -12 gpurun_out/${T}_pytest.txt
