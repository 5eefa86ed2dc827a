"""Timeline of the single-launch batch-norm passes (diagnostics): per-CTA %globaltimer stamps at
entry / loads issued / local sums done / cluster sums done / stores issued, from the instrumented library
(tools/build_trace_lib.sh).   python tools/trace_bn.py [N C H W]"""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from simclr_pytorch_distributed_b200 import _lib   # noqa: E402

_lib.LIB_PATH = os.environ.get('SCB_LIB') or os.path.join(os.path.dirname(os.path.abspath(__file__)), 'bin', 'libtrace.so')
lib = _lib.load()
from simclr_pytorch_distributed_b200 import SyncBatchNorm   # noqa: E402

raw = ctypes.CDLL(_lib.LIB_PATH)
raw.scb_debug_set_bn_trace.argtypes = [ctypes.c_void_p]
raw.scb_debug_set_ring_trace.argtypes = [ctypes.c_void_p]
RING = os.environ.get('SCB_BN_FUSED_MODE', 'coop')[0] == 'c'       # which kernels run (and whose stamps to read)
if RING:
    raw.scb_debug_set_bn_trace = raw.scb_debug_set_ring_trace
shape = tuple(int(a) for a in sys.argv[1:5]) if len(sys.argv) >= 5 else (256, 256, 8, 8)
dev = torch.device('cuda')
pool = 8
xs = [(torch.randn(shape, device=dev) * 1.5 + 0.5).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
      .requires_grad_(True) for _ in range(pool)]
dys = [torch.randn(shape, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last) for _ in range(pool)]
bn = SyncBatchNorm(shape[1]).to(dev)
for i in range(pool):
    bn(xs[i]).backward(dys[i])
torch.cuda.synchronize()
NAMES = (['entry', 'local sums', 'reductions', 'grid barrier', 'table built', 'stores issued'] if RING else
         ['entry', 'loads issued', 'local sums', 'cluster sums', 'stores issued'])


def show(tag, tr):
    t = tr.cpu().view(-1, 8)
    t = t[t[:, 0] > 0][:, :len(NAMES)]
    b = int(t[:, 0].min())
    t = (t - b).float() / 1e3
    print('%s: %d CTAs' % (tag, t.shape[0]))
    for j, n in enumerate(NAMES):
        c = t[:, j]
        print('  %-14s min %6.2f  median %6.2f  max %6.2f us' % (n, c.min(), c.median(), c.max()))


for which in ('fwd', 'bwd'):
    for rep in range(2):
        trace = torch.zeros(256 * 8, dtype=torch.int64, device=dev)
        i = (3 + rep) % pool
        x = xs[i]
        x.grad = None
        if which == 'fwd':
            assert raw.scb_debug_set_bn_trace(trace.data_ptr()) == 0
            y = bn(x)
            torch.cuda.synchronize()
            raw.scb_debug_set_bn_trace(None)
            y.backward(dys[i])
        else:
            y = bn(x)
            torch.cuda.synchronize()
            assert raw.scb_debug_set_bn_trace(trace.data_ptr()) == 0
            y.backward(dys[i])
            torch.cuda.synchronize()
            raw.scb_debug_set_bn_trace(None)
        torch.cuda.synchronize()
        show('%s rep %d' % (which, rep), trace)
