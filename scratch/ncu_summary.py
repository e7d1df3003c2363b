import csv, sys
rows=list(csv.reader(open(sys.argv[1] if len(sys.argv)>1 else 'gpurun_out/full_raw.csv')))
hdr=rows[0]
keys=['gpu__time_duration.sum','launch__grid_size','launch__registers_per_thread','launch__shared_mem_per_block_dynamic','launch__occupancy_limit_registers','launch__occupancy_limit_shared_mem','launch__waves_per_multiprocessor','sm__warps_active.avg.pct_of_peak_sustained_active','dram__bytes_read.sum','dram__bytes_write.sum','dram__throughput.avg.pct_of_peak_sustained_elapsed','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','smsp__inst_executed.sum','sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_tensor_op_dmma.sum','sm__pipe_tensor_op_dmma_cycles_active.avg.pct_of_peak_sustained_active','smsp__cycles_active.avg','sm__cycles_elapsed.avg','smsp__issue_active.avg.pct_of_peak_sustained_active','smsp__thread_inst_executed_per_inst_executed.ratio','lts__t_sector_hit_rate.pct','l1tex__t_sector_hit_rate.pct']
for r in rows[2:]:
    print('-----', r[hdr.index('Kernel Name')])
    for k in keys:
        if k in hdr: print('  %-70s %s %s'%(k, r[hdr.index(k)], rows[1][hdr.index(k)]))
    out=[]
    for i,h in enumerate(hdr):
        if 'issue_stalled' in h and 'per_issue_active.ratio' in h and 'not_issued' not in h:
            try: out.append((float(r[i]),h.replace('smsp__average_warps_issue_stalled_','').replace('_per_issue_active.ratio','')))
            except: pass
    print('  stalls:', ', '.join('%s %.2f'%(h,v) for v,h in sorted(out,reverse=True)[:7]))
