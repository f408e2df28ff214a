#!/bin/bash
# usage: [GPUS=8] tools/gpu/submit.sh <timeout_s> <command string>   -- gpurun, retried while the pod answers busy (rc 3)
TO=$1; shift
G=""; [ -n "$GPUS" ] && G="--gpus $GPUS"
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun $G --timeout $TO -- "$@"
  rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc (attempt $i)"; exit $rc; fi
  sleep 45
done
echo "gave up after 40 busy answers"; exit 3
