#!/bin/bash
# build, run CPU smoke import, then gpurun the given command (always ship a fresh .so)
set -e
cd /root/repo
bash gndnet_b200/csrc/build.sh > /tmp/build.log 2>&1 || { tail -30 /tmp/build.log; exit 1; }
T=${GPU_TIMEOUT:-900}
/usr/local/graft/bin/gpurun --timeout $T -- "$@"
