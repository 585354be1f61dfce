import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
h = rows[0]; v = rows[-1]
want = ['gpu__time_duration.sum', 'smsp__inst_executed.avg', 'sm__mio_pq_write_cycles_active.avg', 'sm__mio_pq_read_cycles_active.avg',
        'sm__warps_active.avg.pct', 'smsp__issue_active.avg.pct', 'smsp__average_warps_issue_stalled', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active', 'sass__inst_executed_local', 'sm__mio_inst_issued.avg', 'launch__registers_per_thread',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__throughput.avg.pct']
for k, x in zip(h, v):
    if any(w in k for w in want) and not k.endswith('peak_sustained_elapsed') or 'mio_pq' in k and 'avg' in k:
        try:
            if float(x.replace(',', '')) == 0: continue
        except ValueError: pass
        print(k, x)
