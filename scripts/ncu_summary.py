#!/usr/bin/env python
"""Summarise an .ncu-rep (raw page + source page stall totals) into a text file for profiles/.
usage: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep profiles/name.txt"""
import csv, io, subprocess, sys
rep, out = sys.argv[1], sys.argv[2]
KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
        'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__inst_executed.sum', 'sm__cycles_elapsed.avg',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'lts__t_bytes.sum', 'l1tex__t_bytes.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio']
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
lines = []
hdr, units = rows[0], rows[1]
for vals in rows[2:]:
    lines.append('== ' + vals[hdr.index('Kernel Name')] if 'Kernel Name' in hdr else '== kernel')
    for h, u, v in zip(hdr, units, vals):
        if h in KEYS:
            lines.append('%-70s %-14s %s' % (h, u, v))
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h = rows[1]; ix = {k: i for i, k in enumerate(h)}
stalls = [k for k in h if k.startswith('stall_') and 'Not Issued' not in k]
tot = dict.fromkeys(stalls, 0); ts = 0; data = []
for r in rows[2:]:
    try: n = int(r[ix['# Samples']])
    except Exception: continue
    ts += n; data.append(r)
    for s in stalls:
        try: tot[s] += int(r[ix[s]])
        except Exception: pass
lines.append('-- warp stall samples (all): total %d' % ts)
for s, v in sorted(tot.items(), key=lambda x: -x[1])[:8]:
    lines.append('   %-28s %9d  %5.1f%%' % (s, v, 100.0 * v / max(ts, 1)))
lines.append('-- top SASS instructions by samples')
for r in sorted(data, key=lambda r: -int(r[ix['# Samples']] or 0))[:25]:
    lines.append('   %8s  exec %10s  %s' % (r[ix['# Samples']], r[ix['Instructions Executed']], r[ix['Source']][:100]))
open(out, 'w').write('\n'.join(lines) + '\n')
print('\n'.join(lines))
