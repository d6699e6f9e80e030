"""Turns the two ncu outputs of a round into the small files kept under profiles/.

  python profiles/summarize_ncu.py launches <launch_list.csv> <out.csv> "<command the list was taken from>"
      launch_list.csv = `ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file ... <command>`
      out.csv         = per kernel: launches, average duration, share of the time of this library's kernels
  python profiles/summarize_ncu.py full <raw.csv> <out.json> [traffic.json]
      raw.csv         = `ncu -i capture.ncu-rep --page raw --csv`
      out.json        = per captured kernel: duration, grid, registers, DRAM bytes, tensor-pipe / DRAM / SM utilisation
      traffic.json    = dram read+write bytes per launch, keyed "<bench kernel kind>:bf16" (read by bench.py's roofline)
"""
import csv
import json
import sys
from collections import OrderedDict

KINDS = {'fused_mlp_kernel': 'fused_fwd_head', 'tc_gemm_kernel': 'dw_gemm', 'optimizer_kernel': 'optimizer',
         'cast_bf16_kernel': 'cast_bf16', 'wgrad_update_kernel': 'dw_gemm_update', 'fit_step_kernel': 'fit_step_small'}


def rows(path):
    with open(path, newline='') as f:
        lines = [l for l in f if not l.startswith('==')]
    return list(csv.DictReader(lines))


def to_bytes(value, unit):
    v = float(value.replace(',', ''))
    return v * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[unit]


def launches(src, dst, command):
    agg = OrderedDict()
    for r in rows(src):
        if r.get('Metric Name') != 'gpu__time_duration.sum':
            continue
        name = r['Kernel Name'].split('(')[0]
        us = float(r['Metric Value'].replace(',', '')) * {'ns': 1e-3, 'us': 1.0, 'ms': 1e3}[r['Metric Unit']]
        agg.setdefault(name, []).append(us)
    ours = {k: v for k, v in agg.items() if 'tsb::' in k}
    total = sum(sum(v) for v in ours.values())
    with open(dst, 'w') as f:
        f.write(f'# ncu launch list, {command}; cold-cache serialised times: compare SHARES\n')
        f.write('kernel,launches,avg_us,share_of_tsb_kernels\n')
        for k, v in sorted(ours.items(), key=lambda kv: -sum(kv[1])):
            f.write(f'{k},{len(v)},{sum(v) / len(v):.2f},{sum(v) / total:.3f}\n')
        others = {k: v for k, v in agg.items() if k not in ours}
        f.write(f'# other kernels in the capture (torch): {sum(len(v) for v in others.values())} launches, '
                f'{sum(sum(v) for v in others.values()):.1f} us\n')


WANT = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__cluster_size', 'launch__registers_per_thread',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__inst_executed.sum', 'lts__t_bytes.sum']


def full(src, dst, traffic_path=None):
    rs = rows(src)
    units = rs[0]          # the raw page carries a units row right under the header
    out, traffic = [], {}
    for r in rs[1:]:
        rec = {'kernel': r['Kernel Name'][:80]}
        for k in WANT:
            if k in r:
                rec[k] = f"{r[k]} {units.get(k, '')}".strip()
        out.append(rec)
        kind = next((v for k, v in KINDS.items() if k in r['Kernel Name']), None)
        if kind and 'dram__bytes_read.sum' in r:
            b = to_bytes(r['dram__bytes_read.sum'], units['dram__bytes_read.sum']) + \
                to_bytes(r['dram__bytes_write.sum'], units['dram__bytes_write.sum'])
            traffic.setdefault(f'{kind}:bf16', []).append(b)
    json.dump(out, open(dst, 'w'), indent=1)
    if traffic_path:
        json.dump({k: sum(v) / len(v) for k, v in traffic.items()}, open(traffic_path, 'w'), indent=1)


if __name__ == '__main__':
    if sys.argv[1] == 'launches':
        launches(sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else '')
    else:
        full(sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else None)
