"""Timeline of CTA 0 of the tensor-core backward kernel (diagnostics): python tools/trace_bwd.py"""
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from simclr_pytorch_distributed_b200 import _lib, losses as L   # noqa: E402

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
trace = torch.zeros(5 * 1024, dtype=torch.int64, device=dev)
lib.scb_debug_set_trace(trace.data_ptr())
L.launch_backward(lib, prob, plan, stats[0], stats[2], g, dk, dev)
torch.cuda.synchronize()
lib.scb_debug_set_trace(None)
t = trace.cpu()[:3072].view(3, 512, 2)
t0 = min(int(t[r, 0, 1]) for r in range(3) if int(t[r, 511, 0]) > 0)
tc = trace.cpu()
ends = [tc[1023], tc[2047], tc[3071]]
print('kernel entry %d, setup done %d, exit %d (relative to first event)' % tuple(int(e) - t0 for e in ends))
names = {8: 'drain wait', 9: 'dK complete', 10: 'dK stored', 1: 'tile start', 2: 'cols staged', 3: 'S ready', 4: 'S in regs', 5: 'P free', 6: 'P written', 7: 'P published'}
for role, label in ((0, 'MMA issuer'), (1, 'epilogue WG0'), (2, 'epilogue WG2 (h=1)')):
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
tags = {0: 'MMA ', 1: 'E-h0', 2: 'E-h1'}
for role in range(3):
    n = int(t[role, 511, 0])
    for i in range(n):
        ev.append((int(t[role, i, 1]) - t0, role, int(t[role, i, 0])))
ev.sort()
mm = {200: 'wait P[0]', 201: 'wait P[1]', 210: 'P[0] here', 211: 'P[1] here', 220: 'MMA2+MMA1 h0 issued', 221: 'MMA2+MMA1 h1 issued'}
print('== merged, ns 10000..16500')
for clk, role, eid in ev:
    if 10000 <= clk <= 16500:
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
