"""Timeline of CTA 0 of the tensor-core backward kernel (diagnostics): python tools/trace_bwd.py"""
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from simclr_pytorch_distributed_b200 import _lib, losses as L   # noqa: E402

_lib.LIB_PATH = os.environ.get('SCB_LIB') or os.path.join(os.path.dirname(os.path.abspath(__file__)), 'bin', 'libtrace.so')   # tools/build_trace_lib.sh
lib = _lib.load()
dev = torch.device('cuda')
B, V, D = 4096, 2, 128
z = F.normalize(torch.randn(B, V, D, device=dev), dim=-1).to(torch.bfloat16)
plan = L.LossPlan(B, V, D, engine=_lib.ENGINE_BF16)
k = torch.empty((plan.n_k, plan.ldk), dtype=plan.k_torch, device=dev)
stats = torch.empty((4, plan.n_k), dtype=torch.float32, device=dev)
loss = torch.empty(2, dtype=torch.float32, device=dev)
g = torch.ones((), dtype=torch.float32, device=dev)
dk = torch.empty((plan.rows_local, plan.ldk), dtype=torch.float32, device=dev)
prob = L._problem(plan, k, None, None, None, 0.5, 0.07)
L.launch_prep(lib, z, plan, False, k, None, None)
L.launch_forward(lib, prob, plan, stats[0], stats[1], stats[2], loss, dev)
for it in range(3):
    L.launch_backward(lib, prob, plan, stats[0], stats[2], g, dk, dev)
torch.cuda.synchronize()
_a = torch.randn(4096, 4096, device=dev)
for _ in range(200):
    _a @ _a
torch.cuda.synchronize()
trace = torch.zeros(6 * 1024, dtype=torch.int64, device=dev)
if os.environ.get('QUIET'):
    trace[4095] = 1          # per-CTA entry / exit stamps only
if os.environ.get('EVENTS'):
    trace[4094] = sum(1 << (int(e) & 31) for e in os.environ['EVENTS'].split(','))
lib.scb_debug_set_trace(trace.data_ptr())
L.launch_backward(lib, prob, plan, stats[0], stats[2], g, dk, dev)
torch.cuda.synchronize()
lib.scb_debug_set_trace(None)
t = trace.cpu()[:3072].view(3, 512, 2)
t0 = min([int(t[r, 0, 1]) for r in range(3) if int(t[r, 511, 0]) > 0] or [0])
tc = trace.cpu()
ends = [tc[1023], tc[2047], tc[3071]]
print('kernel entry %d, setup done %d, exit %d (relative to first event)' % tuple(int(e) - t0 for e in ends))
names = {12: 'loop top', 2: 'cols here', 8: 'drain wait', 9: 'dK complete', 10: 'dK stored', 1: 'tile start', 2: 'cols staged', 3: 'S ready', 4: 'S in regs', 5: 'P free', 6: 'P written', 7: 'P published'}
for role, label in ((0, 'MMA issuer'), (1, 'epilogue WG0'), (2, 'epilogue WG2')):
    n = int(t[role, 511, 0])
    print('==', label, n, 'events')
    prev = None
    for i in range(min(n, 90)):
        eid, clk = int(t[role, i, 0]), int(t[role, i, 1]) - t0
        d = '' if prev is None else '+%d' % (clk - prev)
        prev = clk
        print('  %6d %-6s %s %s' % (clk, d, eid, names.get(eid, '') if role else ''))

# merged view of a steady-state window
ev = []
tags = {0: 'MMA ', 1: 'E-g0', 2: 'E-g2'}
for role in range(3):
    n = int(t[role, 511, 0])
    for i in range(n):
        ev.append((int(t[role, i, 1]) - t0, role, int(t[role, i, 0])))
ev.sort()
mm = {200: 'wait P', 210: 'P here', 220: 'MMA2 issued', 100: 'mma1: wait B', 101: 'mma1: B here', 102: 'mma1 issued'}
print('== merged, ns 10000..14000')
for clk, role, eid in ev:
    if 10000 <= clk <= 14000:
        print('  %6d  %s %s%s' % (clk, tags[role], '          ' * role, mm.get(eid) if role == 0 else names.get(eid, eid)))

# wall-clock entry / exit of every CTA (globaltimer, ns)
w = trace.cpu()[3072:3072 + 1024].view(512, 2)
w = w[w[:, 0] > 0]
b = int(w[:, 0].min())
st, en = (w[:, 0] - b).float() / 1e3, (w[:, 1] - b).float() / 1e3
life = en - st
print('CTAs %d: start %.2f..%.2f us, end %.2f..%.2f us, lifetime min %.2f / median %.2f / max %.2f us' % (
    len(w), st.min(), st.max(), en.min(), en.max(), life.min(), life.median(), life.max()))
print('CTA 0 lifetime %.2f us' % float(life[0]))
order = torch.argsort(life, descending=True)[:6].tolist()
print('longest CTAs:', [(int(i), round(float(life[i]), 2)) for i in order])

lp = trace.cpu()[4096:4096 + 512].view(256, 2)[:len(w)]
ok = lp[:, 0] > 0
if ok.any():
    ls = (lp[ok, 0] - w[ok, 0]).float() / 1e3
    le = (lp[ok, 1] - w[ok, 0]).float() / 1e3
    ex = (w[ok, 1] - lp[ok, 1]).float() / 1e3
    print('tile loop starts %.2f us after entry (median), last tile reached at %.2f us, then %.2f us to exit' % (
        ls.median(), le.median(), ex.median()))

if os.environ.get('EVENTS'):
    for role, label in ((1, 'team 0'), (2, 'team 1')):
        n = int(t[role, 511, 0])
        evs = [(int(t[role, i, 0]), int(t[role, i, 1]) - t0) for i in range(n)]
        ids = sorted(set(e for e, _ in evs))
        for a in ids:
            ta = [c for e, c in evs if e == a]
            d = [y - x for x, y in zip(ta[3:-2], ta[4:-1])]
            if d:
                print('%s: event %d (%s) period median %d ns' % (label, a, names.get(a, ''), sorted(d)[len(d) // 2]))
        if len(ids) == 2:
            a, b = ids
            ta = [c for e, c in evs if e == a]
            tb = [c for e, c in evs if e == b]
            # delay from each a to the next b
            dl = []
            for x in ta[3:-2]:
                nb = [y for y in tb if y >= x]
                if nb:
                    dl.append(nb[0] - x)
            if dl:
                print('%s: %d -> %d median %d ns' % (label, a, b, sorted(dl)[len(dl) // 2]))

# do CTAs whose span crosses a row-block boundary (an extra mid-kernel drain) take longer?
CT, q = (2 * B * 1 + 127) // 128, None
T = ((V * B + 127) // 128) * CT
import math
q = max(math.ceil(T / 148), math.ceil(CT / 15))
cross = torch.tensor([((c * q) // CT) != ((min(T, (c + 1) * q) - 1) // CT) for c in range(len(life))])
if cross.any() and (~cross).any():
    print('CTAs inside one row block: median %.2f us; crossing a boundary: median %.2f us (%d of %d)' % (
        float(life[~cross].median()), float(life[cross].median()), int(cross.sum()), len(life)))
