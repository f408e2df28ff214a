#!/bin/bash
# Two GPUs, final build: the two-GPU tests, C3 weak and C4 strong at N = 2, and both arms launched the driver's way
export NCCL_DEBUG=WARN
python -m pytest tests/test_records_fetch.py tests/test_drop_in.py -m gpu -q -k "multi or two" > gpurun_out/r2zd_pytest.txt 2>&1; tail -2 gpurun_out/r2zd_pytest.txt
for cfg in C3 C4; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --config $cfg --steps 5 --warmup 3 --no-cpu --no-extras > gpurun_out/r2zd_${cfg}_n2.json 2> gpurun_out/r2zd_${cfg}_n2.err
echo "$cfg rc=$? $(python -c "
import json; d=json.loads(open('gpurun_out/r2zd_${cfg}_n2.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['n_gpus'], d['scaling'], d['e2e']['value'])")"
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r2zd_ref_n2.json 2> gpurun_out/r2zd_ref_n2.err; echo "ref rc=$? $(cut -c1-200 gpurun_out/r2zd_ref_n2.json)"
