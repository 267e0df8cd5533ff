#!/bin/bash
# usage: gpuretry.sh <timeout> [--gpus N] <command...>   retries while the pod has no slot (exit 3)
t=$1; shift
g=1
if [ "$1" == "--gpus" ]; then g=$2; shift 2; fi
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --gpus $g --timeout $t -- "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 90
done
exit 3
