#!/bin/bash
# dev helper: submit a gpurun call, retrying while the pod answers busy/transient (exit 3).
# usage: tools/gpurun_retry.sh <outfile> <timeout_s> [--gpus N] -- '<command>'
out=$1; shift; to=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$to" "$@" > "$out" 2>&1
  rc=$?
  if grep -q "status=transient\|status=busy" "$out" || [ $rc -eq 3 ]; then sleep 45; continue; fi
  exit $rc
done
exit 3
