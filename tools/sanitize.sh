#!/bin/bash
# compute-sanitizer passes over the hot path: memcheck and racecheck on smoke() (the whole fused
# path in two precisions, incl. the tcgen05 / cluster / mbarrier conv kernel) and on the voxelizer parity tests
# (self-resetting device counters, shared-memory rank tables).  Run on the GPU box; summaries land in gpurun_out/.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CS=${CS:-/usr/local/cuda/bin/compute-sanitizer}
VOX_K='not maximum_sizes and not camera_1m'
run() {   # name, tool, command...
  local name=$1 tool=$2; shift 2
  echo "== $name ($tool): $*" | tee gpurun_out/sanitizer_${name}_${tool}.log
  timeout 900 $CS --tool $tool --error-exitcode 7 --print-limit 20 "$@" >> gpurun_out/sanitizer_${name}_${tool}.log 2>&1
  echo "exit code $?" | tee -a gpurun_out/sanitizer_${name}_${tool}.log
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|smoke\[" gpurun_out/sanitizer_${name}_${tool}.log | tail -6
}
run smoke memcheck python -c "import __graft_entry__ as g; g.smoke()"
run smoke racecheck python -c "import __graft_entry__ as g; g.smoke()"
run voxelizer memcheck python -m pytest tests/test_voxelize_gpu.py -q -x -k "$VOX_K"
run voxelizer racecheck python -m pytest tests/test_voxelize_gpu.py -q -x -k "$VOX_K and (small or boundary or one_point or empty or single_cell or ragged or nan)"
