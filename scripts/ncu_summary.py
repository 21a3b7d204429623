"""Summarise an .ncu-rep (raw page + source page) into text: key metrics, stall mix, top stalled SASS.
usage: python scripts/ncu_summary.py gpurun_out/x.ncu-rep [ntop]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); hdr, units, data = rows[0], rows[1], rows[2:]
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__cluster_size', 'sm__cycles_elapsed.max',
        'sm__cycles_elapsed.max.per_second', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.sum', 'sm__inst_executed_pipe_fma.sum',
        'sm__inst_executed_pipe_alu.sum', 'sm__inst_executed_pipe_lsu.sum', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__pipe_xu_cycles_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']
for d in data:
    print('=== kernel:', d[hdr.index('Kernel Name')][:100])
    for i, h in enumerate(hdr):
        if h in keys:
            print(f"  {h:80s} {d[i]:>18s} {units[i]}")
    print('  -- warp stall mix (warps per issue-active cycle) --')
    for i, h in enumerate(hdr):
        if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('_per_issue_active.ratio'):
            try: v = float(d[i])
            except ValueError: continue
            if v > 0.15: print(f"  {h.replace('smsp__average_warps_issue_stalled_','').replace('_per_issue_active.ratio',''):30s} {v:8.2f}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
his = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
if his:
    hi = his[0]; hdr = rows[hi]; end = his[1] - 1 if len(his) > 1 else len(rows)
    data = [r for r in rows[hi + 1:end] if len(r) == len(hdr)]
    ci = {h: i for i, h in enumerate(hdr)}
    samp = ci['# Samples']
    tot = sum(int(r[samp] or 0) for r in data if r[samp].isdigit()) or 1
    stall_cols = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
    agg = {h: sum(int(r[ci[h]] or 0) for r in data if r[ci[h]].isdigit()) for h in stall_cols}
    print('  -- sampled stall reasons (all instructions) --')
    for h, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]:
        print(f"  {h:28s} {100 * v / tot:6.1f}%")
    print(f'  -- top {ntop} SASS by samples (of {tot}) --')
    for r in sorted([r for r in data if r[samp].isdigit()], key=lambda r: -int(r[samp]))[:ntop]:
        st = sorted(((h, int(r[ci[h]] or 0)) for h in stall_cols), key=lambda kv: -kv[1])[:2]
        print(f"  {100 * int(r[samp]) / tot:5.1f}%  {r[ci['Source']][:72]:72s} {st[0][0]}:{st[0][1]} {st[1][0]}:{st[1][1]}")
