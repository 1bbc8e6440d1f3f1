#!/bin/bash
# gpurun with retries while the pod answers "busy" (nothing is charged for those).  Usage:
#   tools/gpurun_retry.sh [--gpus N] --timeout S -- 'command'
for i in $(seq 1 ${GPURUN_TRIES:-12}); do
  /usr/local/graft/bin/gpurun "$@"
  st=$(python3 -c "import json;print(json.load(open('/root/repo/gpurun_out/.last_call.json')).get('status'))" 2>/dev/null)
  if [ "$st" != "transient" ]; then exit 0; fi
  echo "[retry $i] pod busy, sleeping 60 s"
  sleep 60
done
