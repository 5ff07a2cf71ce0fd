#!/bin/bash
# Developer helper: run a command on the GPU box through gpurun, retrying while the pod answers "busy / transient"
# (nothing is charged for those).  Usage: tools/gpurun_retry.sh <timeout-seconds> [--gpus N] -- '<command>'
T=$1; shift
for i in $(seq 1 12); do
  out=$(/usr/local/graft/bin/gpurun --timeout "$T" "$@" 2>&1); rc=$?
  if echo "$out" | grep -q "status=transient\|no box or slot\|retry in a few minutes"; then
    echo "[retry $i] transient, sleeping 90 s"; sleep 90; continue
  fi
  echo "$out" | tail -80
  exit $rc
done
echo "gave up after 12 transient answers"; exit 3
