#!/bin/bash
# One 8-GPU box: BASELINE config 4 (2^20 problems, strong scaling over 2/4/8 ranks, NCCL all-gather of the summaries),
# C3 weak scaling with the float32-record e2e and the top-k e2e, the C++ program through b200::multi_gpu_planner,
# and the two-GPU test of scb_multi_*.
T=${1:-r2g}
export NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT
run() { # N, tag, args...
  local N=$1 tag=$2; shift 2
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py --gpus $N "$@" \
      > gpurun_out/${T}_${tag}_n${N}.json 2> gpurun_out/${T}_${tag}_n${N}.err
  echo "$tag N=$N rc=$? $(head -c 300 gpurun_out/${T}_${tag}_n${N}.json | cut -c1-200)"
  grep -E "NCCL INFO (comm|Connected|ncclCommInitRank)" gpurun_out/${T}_${tag}_n${N}.err | grep -E "nranks|Init COMPLETE" | head -3 > gpurun_out/${T}_${tag}_n${N}.nccl.txt
  grep -v "NCCL INFO" gpurun_out/${T}_${tag}_n${N}.err | tail -5
}
python -m pytest tests/test_records_fetch.py tests/test_drop_in.py -m gpu -q -k "multi" > gpurun_out/${T}_pytest_multi.txt 2>&1; tail -3 gpurun_out/${T}_pytest_multi.txt
python bench.py --config C4 --steps 3 --warmup 3 --no-cpu --no-extras > gpurun_out/${T}_c4_n1.json 2> gpurun_out/${T}_c4_n1.err; echo "c4 N=1 rc=$?"
for N in 2 4 8; do run $N c4 --config C4 --steps 5 --warmup 3 --no-cpu --no-extras; done
for N in 2 4 8; do run $N c3 --steps 10 --warmup 3 --no-cpu --no-extras; done
g++ -std=c++20 -O2 -Iinclude tests/cpp/multi_gpu_c4.cpp -o /tmp/multi_gpu_c4 -Lsea-current_b200/csrc -lseacurrent_b200 -Wl,-rpath,$PWD/sea-current_b200/csrc
for N in 1 2 4 8; do /tmp/multi_gpu_c4 $N 1048576 2 > gpurun_out/${T}_cpp_c4_n${N}.json 2>> gpurun_out/${T}_cpp.err; echo "cpp N=$N rc=$? $(cat gpurun_out/${T}_cpp_c4_n${N}.json)"; done
nvidia-smi --query-gpu=index,name,memory.used --format=csv | head -10
