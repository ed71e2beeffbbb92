"""Machine-readable digest of this round's ncu captures of the flux kernel -> profiles/r02_flux_kernel_ncu.json,
which bench.py copies into its JSON line (labelled as coming from profiles/, never measured inside bench.py).

    python tools/ncu_record.py --f64 gpurun_out/flux_full.ncu-rep [--f32 gpurun_out/flux_full_f32.ncu-rep]
                               [--dram gpurun_out/flux_dram_1M.csv --dram-columns 1000000] --out profiles/r02_flux_kernel_ncu.json
"""
import argparse
import csv
import io
import json
import subprocess

PICK = {
    'gpu__time_duration.sum': 'duration',
    'launch__registers_per_thread': 'registers_per_thread',
    'launch__shared_mem_per_block_dynamic': 'dynamic_shared_memory',
    'sm__warps_active.avg.pct_of_peak_sustained_active': 'warps_active_pct',
    'sm__issue_active.avg.pct_of_peak_sustained_elapsed': 'issue_active_pct',
    'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active': 'fp64_pipe_cycles_active_pct',
    'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active': 'fma_pipe_pct',
    'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active': 'alu_pipe_pct',
    'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active': 'lsu_pipe_pct',
    'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active': 'xu_pipe_pct',
    'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed': 'lsu_shared_wavefronts_pct',
    'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum': 'shared_bank_conflicts',
    'lts__throughput.avg.pct_of_peak_sustained_elapsed': 'l2_throughput_pct',
    'lts__t_sector_hit_rate.pct': 'l2_hit_rate_pct',
    'dram__bytes_read.sum': 'dram_bytes_read',
    'dram__bytes_write.sum': 'dram_bytes_write',
    'smsp__inst_executed.sum': 'warp_instructions',
}
SCALE = {'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'byte': 1.0, 'ms': 1e-3, 'us': 1e-6, 'ns': 1e-9, 's': 1.0}


def raw(rep):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    return rows[0], rows[1], rows[2:]


def digest(rep, note):
    hdr, units, rows = raw(rep)
    d = dict(zip(hdr, rows[0]))
    rec = {'from': note, 'kernel': d.get('Kernel Name', '?')[:80]}
    for k, name in PICK.items():
        if k in d:
            u = units[hdr.index(k)]
            v = float(d[k].replace(',', ''))
            rec[name] = v * SCALE.get(u.split('/')[0], 1.0) if u.split('/')[0] in SCALE else v
    stalls = []
    for h, v in d.items():
        if 'pcsamp_warps_issue_stalled' in h and 'not_issued' not in h:
            try:
                stalls.append((float(v.replace(',', '')), h.split('stalled_')[-1]))
            except ValueError:
                pass
    tot = sum(x for x, _ in stalls) or 1.0
    rec['stall_share_pct'] = {h: round(100 * x / tot, 1) for x, h in sorted(stalls, reverse=True)[:8]}
    return rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--f64')
    ap.add_argument('--f32')
    ap.add_argument('--dram', help='csv of an ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum run')
    ap.add_argument('--dram-columns', type=int, default=1000000)
    ap.add_argument('--columns', type=int, default=1480)
    ap.add_argument('--out', required=True)
    a = ap.parse_args()
    rec = {}
    note = 'ncu --set full --clock-control none, tools/profile_flux.py --columns %d, one launch; %s'
    if a.f64:
        rec['f64'] = digest(a.f64, note % (a.columns, a.f64.split('/')[-1]))
    if a.f32:
        rec['f32'] = digest(a.f32, note % (a.columns, a.f32.split('/')[-1]))
    if a.dram:
        rows = [r for r in csv.reader(open(a.dram)) if len(r) > 5]
        hdr = rows[0]
        i_name, i_val, i_unit, i_kern = hdr.index('Metric Name'), hdr.index('Metric Value'), hdr.index('Metric Unit'), hdr.index('Kernel Name')
        dr = {'columns': a.dram_columns,
              'from': 'ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none, flux_kernel launch of '
                      'tools/profile_flux.py --columns %d; %s' % (a.dram_columns, a.dram.split('/')[-1])}
        for r in rows[1:]:
            if 'flux_kernel' not in r[i_kern]:
                continue
            v = float(r[i_val].replace(',', '')) * SCALE.get(r[i_unit], 1.0)
            dr['dram_bytes_read' if 'read' in r[i_name] else 'dram_bytes_write'] = v
        rec['dram_1M_columns'] = dr
    json.dump(rec, open(a.out, 'w'), indent=1)
    print(json.dumps(rec, indent=1))


if __name__ == '__main__':
    main()
