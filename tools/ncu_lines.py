#!/usr/bin/env python
"""Per-source-line instruction counts of one kernel from an ncu report (needs -lineinfo + --import-source on).
Usage: ncu_lines.py rep.ncu-rep kernel_regex [top_n]   (first matching launch only)"""
import csv, io, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--print-source', 'cuda,sass', '--csv', '--kernel-name', 'regex:' + pat],
                     capture_output=True, text=True).stdout
out, fname, hdr, kernels = [], None, None, 0
for r in csv.reader(io.StringIO(raw)):
    if not r:
        continue
    if r[0] == 'Kernel Name':
        kernels += 1
        if kernels > 1:
            break
        continue
    if r[0] == 'File Name':
        fname = r[1].split('/')[-1]; continue
    if r[0] == 'Line No':
        hdr = r; continue
    if hdr and r[0].isdigit():
        iE, iS = hdr.index('Instructions Executed'), hdr.index('# Samples')
        try:
            e, s = int(r[iE] or 0), int(r[iS] or 0)
        except ValueError:
            continue
        if e:
            out.append((e, s, fname, int(r[0]), r[1].strip()[:100]))
tot = sum(o[0] for o in out); ts = sum(o[1] for o in out) or 1
print(f'total warp-instructions {tot}, samples {ts}')
for e, s, f, ln, src in sorted(out, reverse=True)[:top]:
    print(f'{e/tot:6.3f} {s/ts:6.3f} {f}:{ln:4d}  {src}')
