#!/bin/bash
# usage: gpurun_retry.sh <timeout_s> [--gpus N] -- '<command>'   retries while the pod answers busy (nothing is charged for those)
T=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$T" "$@" > /tmp/gpurun_try.log 2>&1
  rc=$?
  if grep -q "status=transient" /tmp/gpurun_try.log || [ $rc -eq 3 ]; then
    echo "[retry $i] busy, sleeping 120 s"; sleep 120; continue
  fi
  cat /tmp/gpurun_try.log; exit $rc
done
echo "gave up after 40 tries"; exit 3
