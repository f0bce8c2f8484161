#!/bin/bash
# Developer aid: run a gpurun command, retrying while the pod answers "busy" (exit code 3: nothing charged).
#   tools/gpurun_retry.sh <timeout-seconds> '<command>'  [extra gpurun args, e.g. --gpus 2]
t=$1; cmd=$2; shift 2
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$t" "$@" -- "$cmd"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  echo "[retry] busy, attempt $i; sleeping 45 s" >&2
  sleep 45
done
exit 3
