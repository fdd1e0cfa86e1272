#!/bin/bash
# gpurun with retries while the pod answers "busy" (exit 3 / transient); args as tools/g.sh
cd /root/repo
for i in $(seq 1 ${GPU_RETRIES:-20}); do
  out=$(bash tools/g.sh "$@" 2>&1); rc=$?
  echo "$out" | tail -${GPU_TAIL:-12}
  if echo "$out" | grep -q "status=transient\|nothing was charged"; then sleep 120; continue; fi
  exit $rc
done
exit 3
