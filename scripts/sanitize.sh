#!/bin/bash
# compute-sanitizer over the kernel test-suite.  usage: scripts/sanitize.sh [memcheck|racecheck|synccheck] [pytest -k expr]
# NOTE: on the pool this was developed on compute-sanitizer is closed (the wrapper refuses to start it; tried again in
# round 2: gpurun_out/r2_sanitizer_memcheck.log / profiles/sanitizer_attempt_r2.log), so this target has not been run there;
# the kernels' own guards are bounded mbarrier / peer-flag waits that trap, shape checks in every launcher, ragged-shape
# tests (T = 300, N = 3352, K = 48) and bit-for-bit rerun comparisons.
# One tool per gpurun call; memcheck slows the kernels ~10-50x, so the default selection is the small-shape tests.
# The summary ("ERROR SUMMARY: N errors") lands in gpurun_out/sanitizer_<tool>.log.
cd "$(dirname "$0")/.."
TOOL=${1:-memcheck}
SEL=${2:-"not 1024 and not model"}
timeout 3400 /usr/local/graft/bin/gpurun --timeout 900 -- "timeout 800 compute-sanitizer --tool $TOOL --error-exitcode 3 --print-limit 20 python -m pytest tests/test_kernels_gpu.py tests/test_ssd_gpu.py tests/test_sscan_gpu.py tests/test_decode_gpu.py -x -q -k '$SEL' > gpurun_out/sanitizer_$TOOL.log 2>&1; echo rc=\$?; grep -E 'ERROR SUMMARY|passed|failed' gpurun_out/sanitizer_$TOOL.log | tail -4"
