#!/usr/bin/env python
"""profiles/r02_ncu_targets.md from the raw ncu CSV of tools/ncu_targets.py:

    ncu -i r02_targets.ncu-rep --page raw --csv > gpurun_out/r02_targets_raw.csv
    python tools/ncu_summary.py gpurun_out/r02_targets_raw.csv gpurun_out/ncu_targets_order.json
"""
import csv
import json
import sys

WANT = [
    ('gpu__time_duration.sum', 'time'),
    ('dram__bytes_read.sum', 'dram read'),
    ('dram__bytes_write.sum', 'dram write'),
    ('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'DRAM %'),
    ('lts__t_sector_hit_rate.pct', 'L2 hit %'),
    ('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'tensor pipe %'),
    ('sm__inst_executed_pipe_tensor_op_gmma.avg.pct_of_peak_sustained_active', 'tensor op %'),
    ('sm__warps_active.avg.pct_of_peak_sustained_active', 'warps active %'),
    ('launch__registers_per_thread', 'regs'),
    ('launch__grid_size', 'grid'),
]


def to_bytes(v, unit):
    v = float(v.replace(',', ''))
    u = unit.lower()
    scale = {'byte': 1, 'kbyte': 1e3, 'mbyte': 1e6, 'gbyte': 1e9, 'tbyte': 1e12}.get(u, 1)
    return v * scale


def to_ms(v, unit):
    v = float(v.replace(',', ''))
    return v * {'ns': 1e-6, 'us': 1e-3, 'usecond': 1e-3, 'ms': 1.0, 'msecond': 1.0, 'second': 1e3,
                's': 1e3, 'nsecond': 1e-6}.get(unit.lower(), 1e-6)


def main():
    rows = list(csv.reader(open(sys.argv[1], newline='')))
    # header row, unit row, then one row per launch
    hdr_i = next(i for i, r in enumerate(rows) if 'Kernel Name' in r)
    hdr, units = rows[hdr_i], rows[hdr_i + 1]
    col = {name: i for i, name in enumerate(hdr)}
    order = json.load(open(sys.argv[2])) if len(sys.argv) > 2 else []
    launches = rows[hdr_i + 2:]
    out = ['| # | launch (tools/ncu_targets.py) | kernel | time ms | DRAM read GB | DRAM write GB | DRAM GB/s | '
           'DRAM % | L2 hit % | tensor pipe % | warps active % | regs | grid |', '|' + '---|' * 13]
    traffic = {}
    k = 0
    for r in launches:
        if len(r) < len(hdr):
            continue
        name = r[col['Kernel Name']]
        short = name.split('(')[0].split('<')[0].split('::')[-1]
        if 'fixup' in name:
            continue

        def get(metric):
            i = col.get(metric)
            return (r[i], units[i]) if i is not None and r[i] != '' else (None, None)
        t, tu = get('gpu__time_duration.sum')
        ms = to_ms(t, tu) if t else float('nan')
        rd, ru = get('dram__bytes_read.sum')
        wr, wu = get('dram__bytes_write.sum')
        rdb = to_bytes(rd, ru) if rd else float('nan')
        wrb = to_bytes(wr, wu) if wr else float('nan')
        label = order[k] if k < len(order) else '?'
        vals = []
        for m in ('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct',
                  'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
                  'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
                  'launch__grid_size'):
            v, _ = get(m)
            vals.append(v if v is not None else '-')
        out.append('| {} | {} | `{}` | {:.3f} | {:.2f} | {:.2f} | {:.0f} | {} |'.format(
            k, label, short, ms, rdb / 1e9, wrb / 1e9, (rdb + wrb) / ms / 1e6 if ms == ms else 0,
            ' | '.join(vals)))
        traffic[label] = {'dram_bytes': rdb + wrb, 'ms_under_ncu': ms, 'kernel': short}
        k += 1
    print('\n'.join(out))
    json.dump(traffic, open('gpurun_out/r02_ncu_traffic.json', 'w'), indent=1)


if __name__ == '__main__':
    main()
