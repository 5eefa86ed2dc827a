"""Summarise an `ncu --page source --csv` dump: top SASS instructions by stall samples and the
totals per stall reason.   python tools/ncu_src_top.py src.csv [topN]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
reasons = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
data = []
tot = {r: 0 for r in reasons}
total = 0
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    if not r[ix["# Samples"]].isdigit():
        continue
    n = int(r[ix["# Samples"]] or 0)
    total += n
    for s in reasons:
        tot[s] += int(r[ix[s]] or 0)
    data.append((n, r))
print('total samples', total)
print('by reason:', ', '.join('%s=%d' % (k[6:], v) for k, v in sorted(tot.items(), key=lambda kv: -kv[1]) if v))
data.sort(key=lambda t: -t[0])
for n, r in data[:top]:
    rs = sorted(((int(r[ix[s]] or 0), s[6:]) for s in reasons), reverse=True)[:3]
    print('%6d %5.1f%%  %-70s %s  exec=%s' % (n, 100.0 * n / max(total, 1), r[ix['Source']].strip()[:70],
                                             ' '.join('%s:%d' % (b, a) for a, b in rs if a), r[ix['Instructions Executed']]))
