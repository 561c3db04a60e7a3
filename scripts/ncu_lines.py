#!/usr/bin/env python
"""Per source line totals (instructions executed, stall samples) of an .ncu-rep captured with
-lineinfo + --import-source on.  usage: python scripts/ncu_lines.py rep [top] [samples]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
key = 3 if (len(sys.argv) > 3 and sys.argv[3] == 'samples') else 2
txt = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr = None; out = []
for r in rows:
    if r and r[0] == 'Line No':
        hdr = r; continue
    if hdr is None or len(r) < 10: continue
    if r[0] == '': continue          # sass rows
    try:
        out.append((int(r[0]), r[1], int(r[7]), int(r[6])))
    except Exception:
        pass
ti = sum(o[2] for o in out); ts = sum(o[3] for o in out)
print('total inst %d samples %d' % (ti, ts))
for ln, src, inst, smp in sorted(out, key=lambda o: -o[key])[:top]:
    print('%5d %6.2f%% inst %6.2f%% smp  %s' % (ln, 100.0 * inst / ti, 100.0 * smp / max(ts, 1), src.strip()[:110]))
