#!/usr/bin/env bash
# Dev helper (build container): gpurun with retries while the pod answers "busy" (nothing charged).
# Usage: tools/grun.sh <tag> [--gpus N] [--timeout S] -- '<command>'      log: gpurun_out/<tag>_call.log
TAG=$1; shift
mkdir -p gpurun_out
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > gpurun_out/${TAG}_call.log 2>&1
  rc=$?
  if grep -q "status=transient\|status=busy" gpurun_out/${TAG}_call.log || [ $rc -eq 3 ]; then sleep 45; continue; fi
  break
done
tail -12 gpurun_out/${TAG}_call.log
exit $rc
