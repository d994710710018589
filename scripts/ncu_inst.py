"""Instruction distribution of one kernel launch by SASS line range: scripts/ncu_inst.py rep regex launch bucket"""
import csv, subprocess, sys, io
rep, rx = sys.argv[1], sys.argv[2]
which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
bucket = int(sys.argv[4]) if len(sys.argv) > 4 else 50
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + rx], capture_output=True, text=True).stdout
blocks, cur = [], None
for row in csv.reader(io.StringIO(out)):
    if row and row[0] == 'Kernel Name':
        cur = {'name': row[1], 'rows': []}; blocks.append(cur)
    elif cur is not None:
        cur['rows'].append(row)
b = blocks[which]; hdr = b['rows'][0]; ci = {h: i for i, h in enumerate(hdr)}; rows = b['rows'][1:]
tot = sum(int(r[ci['Instructions Executed']] or 0) for r in rows)
tots = sum(int(r[ci['# Samples']] or 0) for r in rows)
print(b['name'], 'total warp-inst', tot, 'samples', tots)
for lo in range(0, len(rows), bucket):
    seg = rows[lo:lo + bucket]
    n = sum(int(r[ci['Instructions Executed']] or 0) for r in seg)
    s = sum(int(r[ci['# Samples']] or 0) for r in seg)
    ops = {}
    for r in seg:
        op = r[ci['Source']].strip().split()
        op = [o for o in op if not o.startswith('@')][0].split('.')[0] if op else '?'
        ops[op] = ops.get(op, 0) + int(r[ci['Instructions Executed']] or 0)
    topops = sorted(ops.items(), key=lambda x: -x[1])[:5]
    print(f"{lo:5d}-{lo+len(seg)-1:5d} inst {100.0*n/tot:5.1f}%  samples {100.0*s/max(tots,1):5.1f}%  {[(k, round(100.0*v/tot,1)) for k,v in topops]}")
