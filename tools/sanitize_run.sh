#!/bin/bash
# compute-sanitizer passes over the small GPU paths (smoke = 5xnd, 20 frames, every kernel of the shared-memory path;
# water / large / xtc / dcd / pca tests for the other kernels).  Run on the GPU box: bash tools/sanitize_run.sh
# Output: gpurun_out/sanitize_<tool>_<what>.log, one summary line each on stdout.
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
run() {   # tool, tag, command...
    local tool=$1 tag=$2; shift 2
    local log=gpurun_out/sanitize_${tool}_${tag}.log
    timeout 420 $CS --tool $tool --error-exitcode 86 --print-limit 20 "$@" > $log 2>&1
    local rc=$?
    echo "$tool $tag rc=$rc $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $log | tail -1) $(grep -E 'passed|failed|smoke ok' $log | tail -1)"
}
SMOKE='import __graft_entry__ as g; g.smoke()'
run memcheck smoke python -c "$SMOKE"
run racecheck smoke python -c "$SMOKE"
run synccheck smoke python -c "$SMOKE"
run memcheck water python -m pytest tests/test_gpu_water.py -x -q -m gpu
run memcheck io python -m pytest tests/test_gpu_xtc.py tests/test_gpu_dcd.py tests/test_gpu_pca.py -x -q -m gpu
run memcheck large python -m pytest tests/test_gpu_large.py -x -q -m gpu
run racecheck water python -m pytest tests/test_gpu_water.py -x -q -m gpu
