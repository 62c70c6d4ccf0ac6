#!/bin/bash
# usage: scan_exp.sh "<extra nvcc flags, e.g. -DMB_TBLK=4 -DMB_TBWD=1>"  -- clock64 timeline of one scan CTA: epilogue / scale durations
cd "$(dirname "$0")/.."
MB_NVCC_EXTRA="-DMB_SSD_TIMING $1" python -m mamba_distributed_b200.ops.build -f >/dev/null
timeout 3000 /usr/local/graft/bin/gpurun --timeout 600 -- 'timeout 300 python scripts/ncu_targets.py ssd > gpurun_out/timeline.txt 2>&1' | tail -1
python -m mamba_distributed_b200.ops.build -f >/dev/null
python - <<'PY'
lines=[l.split() for l in open('gpurun_out/timeline.txt') if l.startswith('T ')]
n=len(lines)//3
ev=sorted(((int(w),int(h),c,int(t)) for _,w,h,c,t in lines[-n:]), key=lambda e:e[3])
d={}
for w,h,c,t in ev: d.setdefault((w,h,c),[]).append(t)
print("epi a->e:", [d[(8,h,'e')][0]-d[(8,h,'a')][0] for h in range(6)], "a times", [d[(8,h,'a')][0] for h in range(6)], "end", d[(8,5,'e')][0])
for w in (0,3,7):
    print(f"scale w{w} half durations:", [[hd-sd for sd,hd in zip(d[(w,h,'s')],d[(w,h,'h')])] for h in range(6)])
PY
