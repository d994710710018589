"""Top stall-sample SASS lines of one kernel from an .ncu-rep (source page): scripts/ncu_hot.py rep regex [launch_idx] [top]"""
import csv, subprocess, sys, io
rep, rx = sys.argv[1], sys.argv[2]
which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + rx], capture_output=True, text=True).stdout
# the csv holds one block per launch, separated by a "Kernel Name" row
blocks, cur = [], None
for row in csv.reader(io.StringIO(out)):
    if row and row[0] == 'Kernel Name':
        cur = {'name': row[1], 'rows': []}
        blocks.append(cur)
    elif cur is not None:
        cur['rows'].append(row)
b = blocks[which]
hdr = b['rows'][0]
ci = {h: i for i, h in enumerate(hdr)}
rows = b['rows'][1:]
tot = sum(int(r[ci['# Samples']] or 0) for r in rows)
tot_inst = sum(int(r[ci['Instructions Executed']] or 0) for r in rows)
print(b['name'], 'launch', which, 'of', len(blocks), 'samples', tot, 'warp-inst', tot_inst, 'sass lines', len(rows))
stall_cols = [h for h in hdr if h.startswith('stall_')]
order = sorted(range(len(rows)), key=lambda i: -int(rows[i][ci['# Samples']] or 0))[:top]
for i in sorted(order):
    r = rows[i]
    s = int(r[ci['# Samples']] or 0)
    st = sorted(((int(r[ci[c]] or 0), c[6:]) for c in stall_cols), reverse=True)[:2] if stall_cols else []
    print(f"{i:5d} {100.0*s/max(tot,1):5.1f}% inst={r[ci['Instructions Executed']]:>9s} {r[ci['Source']].strip()[:90]:90s} {st}")
