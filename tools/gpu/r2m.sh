#!/bin/bash
T=${1:-r2m}
python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.txt
tail -5 gpurun_out/${T}_pytest.txt
