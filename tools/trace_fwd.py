"""Prologue / epilogue timeline of CTA 0 of the tensor-core forward kernel (diagnostics).
python tools/trace_fwd.py [B]"""
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from simclr_pytorch_distributed_b200 import _lib, losses as L   # noqa: E402

_lib.LIB_PATH = os.environ.get('SCB_LIB') or os.path.join(os.path.dirname(os.path.abspath(__file__)), 'bin', 'libtrace.so')   # tools/build_trace_lib.sh
lib = _lib.load()
dev = torch.device('cuda')
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
V, D = 2, 128
z = F.normalize(torch.randn(B, V, D, device=dev), dim=-1).to(torch.bfloat16)
plan = L.LossPlan(B, V, D, engine=_lib.ENGINE_BF16)
k = torch.empty((plan.n_k, plan.ldk), dtype=plan.k_torch, device=dev)
stats = torch.empty((4, plan.n_k), dtype=torch.float32, device=dev)
loss = torch.empty(2, dtype=torch.float32, device=dev)
prob = L._problem(plan, k, None, None, None, 0.5, 0.07)
L.launch_prep(lib, z, plan, False, k, None, None)
for _ in range(3):
    L.launch_forward(lib, prob, plan, stats[0], stats[1], stats[2], loss, dev)
torch.cuda.synchronize()
_a = torch.randn(4096, 4096, device=dev)
for _ in range(200):
    _a @ _a
torch.cuda.synchronize()
trace = torch.zeros(5 * 1024, dtype=torch.int64, device=dev)
if os.environ.get('QUIET'):
    trace[4095] = 1          # per-CTA entry / exit stamps only
CTA = int(os.environ.get('TRACE_CTA', '0'))
trace[4094] = CTA            # the CTA whose phases are stamped
lib.scb_debug_set_trace(trace.data_ptr())
L.launch_forward(lib, prob, plan, stats[0], stats[1], stats[2], loss, dev)
torch.cuda.synchronize()
lib.scb_debug_set_trace(None)
t = trace.cpu()[:9].tolist()
tn = trace.cpu()[16:25].tolist()
print('CTA %d' % CTA)
names = ['entry', 'setup done (barriers, TMEM alloc)', 'first A tile landed', 'first B tile landed',
         'first S tile ready', 'tiles of the (last) row block done', 'partials out, ticket taken', 'exit',
         'row block merged (0 = this CTA did not merge)']
for n, v, w in zip(names, t, tn):
    print('%-36s %8d cycles %8d ns' % (n, v - t[0], w - tn[0]))

# wall-clock entry / exit of every CTA (globaltimer, ns)
w = trace.cpu()[3072:3072 + 1024].view(512, 2)
w = w[w[:, 0] > 0]
b = int(w[:, 0].min())
st, en = (w[:, 0] - b).float() / 1e3, (w[:, 1] - b).float() / 1e3
life = en - st
print('CTAs %d: start %.2f..%.2f us, end %.2f..%.2f us, lifetime min %.2f / median %.2f / max %.2f us' % (
    len(w), st.min(), st.max(), en.min(), en.max(), life.min(), life.median(), life.max()))
print('CTA %d lifetime %.2f us' % (CTA, float(life[CTA])))
order = torch.argsort(life, descending=True)[:6].tolist()
print('longest CTAs:', [(int(i), round(float(life[i]), 2)) for i in order])
