#!/usr/bin/env python
"""Condense an `ncu --set full` capture into the text summary committed under profiles/ and into
profiles/ncu_traffic.json (the file bench.py reads `roofline.traffic` from).

  ncu -i gpurun_out/prof.ncu-rep --page raw --csv > gpurun_out/prof_raw.csv       (here, no GPU needed)
  python scripts/ncu_summary.py gpurun_out/prof_raw.csv profiles/r02_x_ncu_full_summary.txt [--traffic]

--traffic rewrites profiles/ncu_traffic.json: per kernel, dram__bytes_read.sum + dram__bytes_write.sum of its LAST
captured launch (bytes), with the summary file named as the source."""
import csv
import json
import os
import sys

KEYS = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
        'launch__shared_mem_per_block_dynamic', 'launch__shared_mem_per_block_static', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'launch__waves_per_multiprocessor',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'smsp__inst_executed.sum', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_tensor_op_dmma.sum', 'sm__pipe_tensor_op_dmma_cycles_active.avg.pct_of_peak_sustained_active',
        'smsp__cycles_active.avg', 'sm__cycles_elapsed.avg', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct']
SCALE = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}


def main():
    src = sys.argv[1]
    out_path = sys.argv[2] if len(sys.argv) > 2 and not sys.argv[2].startswith('--') else None
    rows = [r for r in csv.reader(open(src)) if r]
    start = next(i for i, r in enumerate(rows) if 'Kernel Name' in r)
    hdr, units = rows[start], rows[start + 1]
    lines, traffic = [], {}
    for r in rows[start + 2:]:
        name = r[hdr.index('Kernel Name')]
        short = name.split('(')[0].split('::')[-1].replace('void ', '').split('<')[0].strip()
        lines.append('----- ' + name)
        for k in KEYS:
            if k in hdr:
                lines.append('  %-70s %s %s' % (k, r[hdr.index(k)], units[hdr.index(k)]))
        st = []
        for i, h in enumerate(hdr):
            if 'issue_stalled' in h and 'per_issue_active.ratio' in h and 'not_issued' not in h:
                try:
                    st.append((float(r[i]), h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')))
                except ValueError:
                    pass
        lines.append('  stalls: ' + ', '.join('%s %.2f' % (h, v) for v, h in sorted(st, reverse=True)[:7]))
        try:
            rd = float(r[hdr.index('dram__bytes_read.sum')]) * SCALE.get(units[hdr.index('dram__bytes_read.sum')], 1.0)
            wr = float(r[hdr.index('dram__bytes_write.sum')]) * SCALE.get(units[hdr.index('dram__bytes_write.sum')], 1.0)
            traffic[short] = {"dram_bytes": int(rd + wr), "dram_bytes_read": int(rd), "dram_bytes_write": int(wr),
                              "us": float(r[hdr.index('gpu__time_duration.sum')]) * {'ns': 1e-3, 'us': 1.0, 'ms': 1e3}.get(units[hdr.index('gpu__time_duration.sum')], 1.0)}
        except (ValueError, KeyError):
            pass
    text = '\n'.join(lines) + '\n'
    if out_path:
        open(out_path, 'w').write(text)
    else:
        sys.stdout.write(text)
    if '--traffic' in sys.argv:
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        p = os.path.join(root, 'profiles', 'ncu_traffic.json')
        old = json.load(open(p)) if os.path.exists(p) else {"kernels": {}}
        old["kernels"].update(traffic)
        old["source"] = os.path.relpath(out_path, root) if out_path else src
        json.dump(old, open(p, 'w'), indent=1, sort_keys=True)


if __name__ == '__main__':
    main()
