"""Per source line: instructions executed and stall samples of one kernel of an .ncu-rep (needs -lineinfo and --import-source on).
usage: python scripts/ncu_lines.py <rep> <kernel regex> [top N]"""
import csv, subprocess, sys, collections, io
rep, kre = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass', '-k', 'regex:' + kre],
                     capture_output=True, text=True).stdout
cur_file, rows, hdr = None, [], None
agg = collections.defaultdict(lambda: [0, 0, ''])
launches = 0
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] == 'File Path':
        cur_file = r[1].split('/')[-1]; continue
    if r[0] == 'Function Name':
        launches += 1
        if launches > 1 and cur_file is None: pass
        continue
    if r[0] == 'Line No':
        hdr = {h: i for i, h in enumerate(r)}; continue
    if hdr is None or len(r) < len(hdr) - 2:
        continue
    try:
        ln = r[hdr['Line No']]; inst = int(r[hdr['Instructions Executed']] or 0); smp = int(r[hdr['# Samples']] or 0)
    except (ValueError, KeyError):
        continue
    if not ln.strip():
        continue
    k = (cur_file, int(ln))
    agg[k][0] += inst; agg[k][1] += smp
    src = r[1]
    if src and not agg[k][2]: agg[k][2] = src.strip()[:110]
tot_i = sum(v[0] for v in agg.values()) or 1
tot_s = sum(v[1] for v in agg.values()) or 1
print(f"total warp instructions {tot_i}, samples {tot_s}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{k[0]:16s}:{k[1]:5d}  inst {100*v[0]/tot_i:5.1f}%  stall {100*v[1]/tot_s:5.1f}%  {v[2]}")
