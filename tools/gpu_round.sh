#!/bin/bash
# One GPU-box visit: GPU tests (all, not -x), then ncu launch lists (time + dram bytes) at batch 64 and batch 1.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TAG=${1:-r2}
python -m pytest tests -m gpu -q > gpurun_out/${TAG}_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_tests.log
tail -4 gpurun_out/${TAG}_tests.log
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active
python tools/profile_step.py --batch 64 --steps 2 > gpurun_out/${TAG}_ps64.log 2>&1 && \
ncu --metrics $M --clock-control none --launch-skip 17 -c 40 --csv --log-file gpurun_out/${TAG}_launches_b64.csv python tools/profile_step.py --batch 64 --steps 3 > gpurun_out/${TAG}_ncu64.log 2>&1
ncu --metrics $M --clock-control none --launch-skip 17 -c 40 --csv --log-file gpurun_out/${TAG}_launches_b1.csv python tools/profile_step.py --batch 1 --steps 3 > gpurun_out/${TAG}_ncu1.log 2>&1
echo done
