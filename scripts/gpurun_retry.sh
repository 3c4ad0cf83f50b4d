#!/bin/bash
# Retry a gpurun call while the pod answers "busy" (exit 3 / status=transient; nothing is charged for those).
# Usage: bash scripts/gpurun_retry.sh <timeout seconds> '<command>' [extra gpurun flags]
T=$1; CMD=$2; shift 2
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun "$@" --timeout "$T" -- "$CMD" > /tmp/gpurun_last.log 2>&1
  if ! grep -q "status=transient" /tmp/gpurun_last.log; then break; fi
  sleep 60
done
cat /tmp/gpurun_last.log
