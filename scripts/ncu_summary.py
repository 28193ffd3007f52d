"""Print the key metrics of an `ncu --page raw --csv` export (one block per profiled launch)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
want = ['Kernel Name', 'gpu__time_duration.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_tensor.sum',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_bytes.sum', 'lts__t_sector_hit_rate.pct', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'l1tex__data_pipe_lsu_wavefronts.sum', 'sm__cycles_elapsed.max',
        'sm__cycles_active.avg', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__cycles_active.avg', 'sm__inst_executed.sum',
        'launch__grid_size', 'launch__block_size', 'smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio',
        'sm__clock_frequency' if False else 'gpc__cycles_elapsed.avg.per_second']
extra = []
for r in rows[2:]:
    print('-----')
    for w in want + extra:
        for i, h in enumerate(hdr):
            if h == w:
                print(f"{w:78s} {r[i][:60]:>24s} {units[i]}")
