"""Stall-reason shares of the hot loops of a kernel from an `ncu --set full --import-source on` capture.

    python tools/ncu_stalls.py gpurun_out/flux_full_f32.ncu-rep [--top 25] > profiles/r02_..._stalls.txt

A "hot loop" is a maximal run of SASS instructions executed at least half as often as the most executed one.
For every loop: its share of the kernel's warp samples, instructions per iteration, the stall reasons of its
samples, and the instructions with the most samples.
"""
import argparse
import csv
import io
import subprocess


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('rep')
    ap.add_argument('--top', type=int, default=20)
    a = ap.parse_args()
    out = subprocess.run(['ncu', '-i', a.rep, '--page', 'source', '--csv', '--print-source', 'sass'],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    print(rows[0][1] if len(rows[0]) > 1 else rows[0][0])
    hdr, data = rows[1], rows[2:]
    i_src, i_n, i_ex = hdr.index('Source'), hdr.index('# Samples'), hdr.index('Instructions Executed')
    stalls = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
    total = sum(int(r[i_n]) for r in data)
    top_ex = max(int(r[i_ex]) for r in data)
    print('warp samples %d, SASS instructions %d, executed %d' % (total, len(data), sum(int(r[i_ex]) for r in data)))

    def shares(sel):
        n = sum(int(r[i_n]) for r in sel) or 1
        agg = {}
        for r in sel:
            for i in stalls:
                agg[hdr[i][6:]] = agg.get(hdr[i][6:], 0) + int(r[i])
        return n, ', '.join('%s %.1f' % (k, 100.0 * v / n) for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8] if v)

    n, txt = shares(data)
    print('whole kernel: %s' % txt)
    # hot loops: runs of instructions executed >= top_ex / 2, bridging the untaken side of branches (short cold gaps)
    hot = [int(r[i_ex]) >= top_ex / 2 for r in data]
    loops, start, gap = [], None, 0
    for k, h in enumerate(hot):
        if h:
            if start is None:
                start = k
            gap = 0
            end = k
        elif start is not None:
            gap += 1
            if gap > 120:
                loops.append((start, end))
                start = None
    if start is not None:
        loops.append((start, end))
    for lo, hi in loops:
        sel = [r for r in data[lo:hi + 1] if int(r[i_ex]) >= top_ex / 2]
        n, txt = shares(sel)
        iters = max(int(r[i_ex]) for r in sel)
        print('\nloop at SASS lines %d-%d: %.1f %% of the samples, %d hot instructions (the loops of both directions when they are adjacent), %d warp-iterations'
              % (lo, hi, 100.0 * n / total, len(sel), iters))
        print('  stalls (%% of the loop\'s samples): %s' % txt)
        for r in sorted(sel, key=lambda r: -int(r[i_n]))[:a.top]:
            st = sorted(((hdr[i][6:], int(r[i])) for i in stalls if int(r[i])), key=lambda x: -x[1])[:2]
            print('  %6d  %-64s %s' % (int(r[i_n]), r[i_src].strip()[:64], ', '.join('%s %d' % x for x in st)))


if __name__ == '__main__':
    main()
