#!/bin/bash
# Final build on one 8-GPU box, trimmed: C4 (2^20 problems, strong) at N = 1 and 8, C3 (weak) at N = 8, the C++ program at N = 8.
T=${1:-r2zo}
export NCCL_DEBUG=WARN
run() { local N=$1 tag=$2; shift 2
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + N)) bench.py --gpus $N "$@" \
      > gpurun_out/${T}_${tag}_n${N}.json 2> gpurun_out/${T}_${tag}_n${N}.err
  echo "$tag N=$N rc=$? $(python -c "
import json; d=json.loads(open('gpurun_out/${T}_${tag}_n${N}.json').read().strip().splitlines()[-1]); print(round(d['value']), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), 'topk', round(d.get('e2e_topk',{}).get('value',0)))")"; }
python bench.py --config C4 --steps 3 --warmup 3 --no-cpu > gpurun_out/${T}_c4_n1.json 2> gpurun_out/${T}_c4_n1.err; echo "c4 N=1 rc=$? $(python -c "
import json; d=json.loads(open('gpurun_out/${T}_c4_n1.json').read().strip().splitlines()[-1]); print(round(d['value']), round(d['ms_per_step'],3))")"
run 8 c4 --config C4 --steps 5 --warmup 3 --no-cpu
run 8 c3 --steps 10 --warmup 3 --no-cpu
g++ -std=c++20 -O2 -Iinclude tests/cpp/multi_gpu_c4.cpp -o /tmp/multi_gpu_c4 -Lsea-current_b200/csrc -lseacurrent_b200 -Wl,-rpath,$PWD/sea-current_b200/csrc
/tmp/multi_gpu_c4 8 1048576 2 > gpurun_out/${T}_cpp_c4_n8.json 2>> gpurun_out/${T}_cpp.err; echo "cpp N=8 rc=$? $(cat gpurun_out/${T}_cpp_c4_n8.json | cut -c1-300)"
