#!/bin/bash
# build, then gpurun the given command (always ship a fresh .so).  GPU_TIMEOUT seconds, GPUS=N for a multi-GPU box.
set -e
cd /root/repo
bash gndnet_b200/csrc/build.sh > /tmp/build.log 2>&1 || { tail -30 /tmp/build.log; exit 1; }
T=${GPU_TIMEOUT:-900}
if [ -n "${GPUS:-}" ]; then
  /usr/local/graft/bin/gpurun --gpus $GPUS --timeout $T -- "$@"
else
  /usr/local/graft/bin/gpurun --timeout $T -- "$@"
fi
