#!/bin/bash
# clock64 timeline of one ssd_scan_kernel CTA (see MB_SSD_TIMING in csrc/ssd_fwd_tc.cu).
# Builds an instrumented library, runs the SSD kernels once on a B200 via gpurun, restores the normal build.
set -e
cd "$(dirname "$0")/.."
MB_NVCC_EXTRA=-DMB_SSD_TIMING python -m mamba_distributed_b200.ops.build -f >/dev/null
timeout 3000 /usr/local/graft/bin/gpurun --timeout 600 -- 'timeout 300 python scripts/ncu_targets.py ssd > gpurun_out/timeline.txt 2>&1' | tail -2
python -m mamba_distributed_b200.ops.build -f >/dev/null
python - <<'PY'
lines=[l.split() for l in open('gpurun_out/timeline.txt') if l.startswith('T ')]
n=len(lines)//3
ev=sorted(((int(w),int(h),c,int(t)) for _,w,h,c,t in lines[-n:]), key=lambda e:e[3])
for w,h,c,t in ev:
    if w in (0,3,7,8,16): print(f"w{w:2d} h{h} {c} {t}")
PY
