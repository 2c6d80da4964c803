"""Summarise an .ncu-rep: key metrics, stall breakdown, top source lines.  usage: ncu_summary.py rep [n_lines]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; nl = int(sys.argv[2]) if len(sys.argv) > 2 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); hdr, units = rows[0], rows[1]
want = ['Kernel Name', 'Grid Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__waves_per_multiprocessor', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed.sum', 'l1tex__t_sector_hit_rate.pct',
        'lts__t_sector_hit_rate.pct', 'sm__cycles_active.avg', 'sm__cycles_elapsed.max',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__inst_executed_op_shared_atom.sum']
for r in rows[2:]:
    for w in want:
        if w in hdr: print(f"  {w} = {r[hdr.index(w)]} {units[hdr.index(w)]}")
    for i, h in enumerate(hdr):
        if 'issue_stalled' in h and h.endswith('.ratio') and 'not_issued' not in h and float(r[i] or 0) > 0.05:
            print(f"    stall {h.split('issue_stalled_')[1]:45s} {r[i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h = next(i for i, r in enumerate(rows[:10]) if 'Source' in r)
lines, tot = [], 0
for r in rows[h + 1:]:
    if len(r) > 4 and r[0] != "":
        try: s = int(r[4])
        except ValueError: continue
        lines.append((s, r[0], r[1][:105])); tot += s
print('total samples', tot)
for s, l, t in sorted(lines, reverse=True)[:nl]:
    print(f"{s:7d} {100 * s / tot:5.1f}%  L{l}: {t}")
