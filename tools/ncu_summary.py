"""Compact per-kernel summary of an `ncu --page raw --csv` dump (durations, tensor-pipe share, DRAM/L2 bytes, top stalls)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
H = {h: i for i, h in enumerate(hdr)}
keys = ['gpu__time_duration.sum', 'sm__cycles_active.avg', 'sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu_realtime.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_sector_hit_rate.pct', 'smsp__inst_executed.sum', 'launch__registers_per_thread', 'smsp__cycles_active.avg']
for d in data:
    print('----', d[H['Kernel Name']][:70], d[H['Grid Size']])
    for k in keys:
        for h in hdr:
            if h.endswith(k):
                print(f"   {h[-72:]:72s} {d[H[h]]:>16s} {units[H[h]]}")
    st = [(float(d[i].replace(',', '')), h) for i, h in enumerate(hdr)
          if 'issue_stalled' in h and h.endswith('_per_warp_active.pct') and d[i]]
    for v, h in sorted(st, reverse=True)[:7]:
        print(f"      stall {h.split('issue_stalled_')[1][:50]:50s} {v:.2f}")
