"""Turn ncu reports brought back in gpurun_out/ into the tracked summaries under profiles/.

  python tools/make_profile_summary.py <round-tag> <report.ncu-rep> [<report2.ncu-rep> ...]

Writes profiles/<tag>_<report>.md (selected raw metrics per kernel) and merges the per-kernel
DRAM traffic into profiles/traffic.json, which bench.py reads for the `roofline.traffic` field."""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ['gpu__time_duration.sum', 'sm__cycles_elapsed.max', 'sm__cycles_active.avg', 'dram__bytes_read.sum',
        'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'smsp__inst_executed.sum', 'lts__t_bytes.sum']


def to_bytes(val, unit):
    v = float(val.replace(',', ''))
    return v * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(unit, 1)


def main():
    tag, reports = sys.argv[1], sys.argv[2:]
    traffic_path = os.path.join(ROOT, 'profiles', 'traffic.json')
    traffic = json.load(open(traffic_path)) if os.path.exists(traffic_path) else {}
    for rep in reports:
        out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
        rows = list(csv.reader(out.splitlines()))
        hdr, units = rows[0], rows[1]
        name = os.path.splitext(os.path.basename(rep))[0]
        lines = ['# %s -- ncu --set full --clock-control none (cold caches, serialised launches)' % name, '']
        for r in rows[2:]:
            d = dict(zip(hdr, r))
            kname = d.get('Kernel Name', '?')
            short = kname.split('(')[0].replace('void ', '').strip()
            lines.append('## %s' % kname[:110])
            for k in KEYS:
                if k in d:
                    lines.append('- `%s` = %s %s' % (k, d[k], units[hdr.index(k)]))
            lines.append('')
            try:
                rd = to_bytes(d['dram__bytes_read.sum'], units[hdr.index('dram__bytes_read.sum')])
                wr = to_bytes(d['dram__bytes_write.sum'], units[hdr.index('dram__bytes_write.sum')])
                traffic[short] = {'dram_bytes': rd + wr, 'round': tag, 'report': name}
            except (KeyError, ValueError):
                pass
        with open(os.path.join(ROOT, 'profiles', '%s_%s.md' % (tag, name)), 'w') as f:
            f.write('\n'.join(lines))
    with open(traffic_path, 'w') as f:
        json.dump(traffic, f, indent=1, sort_keys=True)
    print('wrote', len(reports), 'summaries; traffic keys:', sorted(traffic))


if __name__ == '__main__':
    main()
