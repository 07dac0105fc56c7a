#!/bin/bash
# Retries a gpurun call while the pod answers "no box / transient" (exit 3). Usage: tools/gpu_retry.sh <timeout_s> [--gpus N] -- <command>
T=$1; shift
EXTRA=()
while [ "$1" != "--" ]; do EXTRA+=("$1"); shift; done
shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$T" "${EXTRA[@]}" -- "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  echo "[gpu_retry] attempt $i: no box, sleeping 90 s" >&2
  sleep 90
done
exit 3
