"""Summarise an ncu report: headline metrics + top stalled SASS lines.  usage: ncu_top.py file.ncu-rep"""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, vals = rows[0], rows[-1]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'launch__grid_size', 'launch__block_size', 'lts__t_sector_hit_rate.pct',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']
for i, h in enumerate(hdr):
    if h in want:
        print(f"{h} = {vals[i]} {rows[1][i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 24
# the source page repeats "<kernel line>, <header>, <rows...>" for every kernel of the report
heads = [i for i, r in enumerate(rows) if '# Samples' in r and 'Source' in r]
for k, h0 in enumerate(heads):
    hdr = rows[h0]
    end = heads[k + 1] - 1 if k + 1 < len(heads) else len(rows)
    iS, isrc, iex = hdr.index('# Samples'), hdr.index('Source'), hdr.index('Instructions Executed')
    data = [r for r in rows[h0 + 1:end] if len(r) == len(hdr) and r[iS].strip().isdigit()]
    sc = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
    title = " ".join(rows[h0 - 1])[:120] if h0 > 0 else ""
    print(f"==== kernel {k}: {title}")
    print("total samples", sum(int(r[iS]) for r in data), "warp-instr", sum(int(r[iex]) for r in data if r[iex].strip().isdigit()))
    agg = {hdr[i]: sum(int(r[i]) for r in data if r[i].strip().isdigit()) for i in sc}
    print(sorted(agg.items(), key=lambda kv: -kv[1])[:8])
    top = sorted(range(len(data)), key=lambda i: -int(data[i][iS]))[:n]
    for i in sorted(top):
        r = data[i]
        rs = sorted([(hdr[c], int(r[c])) for c in sc if r[c].strip().isdigit() and int(r[c]) > 0], key=lambda kv: -kv[1])[:2]
        print(i, r[iS], r[iex], r[isrc].strip()[:64], rs)
