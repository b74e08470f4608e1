#!/bin/bash
# usage: gpu_retry.sh <timeout> <gpus> <command...>; retries while gpurun answers "busy" (exit 3) or "another call running" (exit 2)
T=$1; G=$2; shift 2
for i in $(seq 1 60); do
  if [ "$G" = "1" ]; then /usr/local/graft/bin/gpurun --timeout $T -- "$@"; else /usr/local/graft/bin/gpurun --gpus $G --timeout $T -- "$@"; fi
  rc=$?
  if [ $rc -ne 3 ] && [ $rc -ne 2 ]; then exit $rc; fi
  sleep 60
done
exit 3
