"""Compact text summary of an .ncu-rep (raw page + hottest source lines), for profiles/.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep [n_source_lines]
"""
import csv
import io
import subprocess
import sys

KEYS = [
    'gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
    'launch__shared_mem_per_block_dynamic', 'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers',
    'sm__warps_active.avg.pct_of_peak_sustained_active', 'sm__issue_active.avg.pct_of_peak_sustained_elapsed',
    'smsp__inst_executed.sum', 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
    'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_cbu.avg.pct_of_peak_sustained_active',
    'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
    'lts__t_bytes.sum', 'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
    'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
    'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
    'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
    'smsp__cycles_active.avg', 'sm__cycles_elapsed.max',
]


def page(rep, name):
    extra = []
    out = subprocess.run(['ncu', '-i', rep, '--page', name, '--csv'] + extra, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    if name == 'source':        # drop the leading "Kernel Name" record
        while rows and 'Source' not in rows[0]:
            rows.pop(0)
    return rows


def main():
    rep = sys.argv[1]
    n_src = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    rows = page(rep, 'raw')
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        print('== kernel:', d.get('Kernel Name', '?')[:100])
        for k in KEYS:
            if k in d:
                print('  %-82s %s %s' % (k, d[k], units[hdr.index(k)]))
        stalls = []
        for h, v in d.items():
            if 'pcsamp_warps_issue_stalled' in h and 'not_issued' not in h:
                try:
                    stalls.append((float(v.replace(',', '')), h.split('stalled_')[-1]))
                except ValueError:
                    pass
        tot = sum(x for x, _ in stalls) or 1.0
        print('  stall samples:', ', '.join('%s %.1f%%' % (h, 100 * x / tot) for x, h in sorted(stalls, reverse=True)[:8]))
    src = page(rep, 'source')
    if len(src) > 2:
        h = src[0]
        try:
            i_samp = next(i for i, x in enumerate(h) if x.startswith('# Samples') or x == 'Warp Stall Sampling (All Samples)' or 'Samples' in x)
        except StopIteration:
            return
        i_src = h.index('Source') if 'Source' in h else 1
        i_inst = next((i for i, x in enumerate(h) if x.startswith('Instructions Executed')), None)
        body = []
        for r in src[1:]:
            try:
                body.append((float(r[i_samp].replace(',', '')), r))
            except (ValueError, IndexError):
                pass
        tot = sum(x for x, _ in body) or 1.0
        print('== hottest lines (%s; total %.0f)' % (h[i_samp], tot))
        for x, r in sorted(body, key=lambda t: -t[0])[:n_src]:
            print('  %5.1f%%  inst=%s  %s' % (100 * x / tot, r[i_inst] if i_inst is not None else '?', r[i_src][:110]))


if __name__ == '__main__':
    main()
