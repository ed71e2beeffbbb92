"""Executed-instruction mix per unit of work from the SASS source page of an .ncu-rep.
    python tools/ncu_opmix.py rep.ncu-rep <units (warp-level)> [detail]"""
import csv, sys, io, subprocess, collections
rep = sys.argv[1]; units = float(sys.argv[2]); detail = len(sys.argv) > 3
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = rows[1]; iS = h.index('Source'); iI = h.index('Instructions Executed'); iN = h.index('# Samples')
ops = collections.Counter(); samp = collections.Counter(); tot = 0
for r in rows[2:]:
    try: n = int(r[iI]); s = int(r[iN])
    except Exception: continue
    t = r[iS].split()
    op = t[1] if t[0].startswith('@') else t[0]
    if not detail: op = op.split('.')[0]
    ops[op] += n; samp[op] += s; tot += n
print('total warp-instructions %d = %.1f per warp-unit' % (tot, tot / units))
ssum = sum(samp.values()) or 1
for op, n in ops.most_common(26): print('  %-22s %6.2f /unit   %5.1f%% of samples' % (op, n / units, 100 * samp[op] / ssum))
