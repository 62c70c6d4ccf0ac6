"""Counts the SASS mnemonics that prove which hardware paths each kernel family uses
(tcgen05 = UTCHMMA / UTCBAR / LDTM, TMA = UTMALDG / UTMASTG / UTMAREDG, packed fp32 = FFMA2 / FMUL2 / FADD2,
cp.async = LDGSTS, peer loads = LDG...SYS).  usage: python scripts/sass_evidence.py > profiles/sass_evidence.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "mamba_distributed_b200", "lib", "libmamba_b200.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
fams = ["gemm_bf16_kernel", "ssd_scan_kernel", "ssd_dcb", "ssd_dbdc", "ssd_outer", "conv1d_fwd", "conv1d_bwd",
        "gated_rmsnorm", "add_rmsnorm", "ce_slab", "sscan", "ssd_state_update", "comm_", "adamw"]
keep = re.compile(r"^(UTC|UTMA|LDTM|STTM|LDGSTS|SYNCS|FFMA2|FMUL2|FADD2|MUFU|LDG\.E\.128|STG\.E\.128|ATOMG|REDG|RED\.|UCGABAR|LDG\.E\.\d*\.?STRONG\.SYS|LDG.*SYS)")
cur, counts = None, collections.defaultdict(collections.Counter)
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = next((f for f in fams if f in m.group(1)), None)
        continue
    if cur is None:
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if m and keep.match(m.group(1)):
        counts[cur][m.group(1)] += 1
for f in fams:
    if f in counts:
        print("==", f)
        print(" ".join(f"{k}:{v}" for k, v in sorted(counts[f].items())))
