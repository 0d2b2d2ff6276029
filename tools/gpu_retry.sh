#!/bin/bash
# usage: gpu.sh TIMEOUT [--gpus N] -- 'command'   : retries while the pod is busy (rc 3)
T=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $T "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  echo "[gpu.sh] busy, retry $i"; sleep 90
done
exit 3
