#!/bin/bash
# compute-sanitizer over the kernel test-suite.  usage: scripts/sanitize.sh [memcheck|racecheck|synccheck] [pytest -k expr]
# One tool per gpurun call; memcheck slows the kernels ~10-50x, so the default selection is the small-shape tests.
# The summary ("ERROR SUMMARY: N errors") lands in gpurun_out/sanitizer_<tool>.log.
cd "$(dirname "$0")/.."
TOOL=${1:-memcheck}
SEL=${2:-"not 1024 and not model"}
timeout 3400 /usr/local/graft/bin/gpurun --timeout 900 -- "timeout 800 compute-sanitizer --tool $TOOL --error-exitcode 3 --print-limit 20 python -m pytest tests/test_kernels_gpu.py tests/test_ssd_gpu.py tests/test_sscan_gpu.py tests/test_decode_gpu.py -x -q -k '$SEL' > gpurun_out/sanitizer_$TOOL.log 2>&1; echo rc=\$?; grep -E 'ERROR SUMMARY|passed|failed' gpurun_out/sanitizer_$TOOL.log | tail -4"
