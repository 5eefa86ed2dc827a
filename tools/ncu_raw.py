"""Print selected metrics per kernel from an ncu report: python tools/ncu_raw.py report.ncu-rep [metric-substr ...]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
extra = sys.argv[2:]
out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
keys = ['gpu__time_duration.sum', 'sm__cycles_elapsed.max', 'sm__cycles_active.avg',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'launch__occupancy_limit_registers',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_bytes.sum',
        'launch__waves_per_multiprocessor', 'smsp__inst_executed.sum']
for r in rows[2:]:
    d = dict(zip(hdr, r))
    print('----', d.get('Kernel Name', '')[:90], '| grid', d.get('launch__grid_size'), 'block', d.get('launch__block_size'))
    for k in keys[:10] + keys[13:]:
        if k in d:
            print('   %-70s %s %s' % (k, d[k], units[hdr.index(k)]))
    for e in extra:
        for h in hdr:
            if e in h:
                print('   %-70s %s %s' % (h, d[h], units[hdr.index(h)]))
