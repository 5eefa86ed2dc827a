"""Host-side cost of one SyncBatchNorm forward+backward through the module (cProfile), NHWC and NCHW."""
import cProfile
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from simclr_pytorch_distributed_b200 import SyncBatchNorm   # noqa: E402

dev = torch.device('cuda')
for cl in (True, False):
    x = torch.randn(64, 64, 8, 8, device=dev, dtype=torch.bfloat16)
    if cl:
        x = x.contiguous(memory_format=torch.channels_last)
    x.requires_grad_(True)
    dy = torch.randn_like(x)
    bn = SyncBatchNorm(64).to(dev)
    ref = torch.nn.BatchNorm2d(64).to(dev)

    def step(m):
        x.grad = None
        m.weight.grad = None
        m.bias.grad = None
        m(x).backward(dy)

    for m, name in ((bn, 'ours'), (ref, 'torch.nn.BatchNorm2d')):
        for _ in range(20):
            step(m)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(300):
            step(m)
        torch.cuda.synchronize()
        print('%s %s: %.1f us per fwd+bwd (host-bound, tiny tensor)' % ('NHWC' if cl else 'NCHW', name, (time.perf_counter() - t0) / 300 * 1e6))
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(300):
        step(bn)
    pr.disable()
    torch.cuda.synchronize()
    st = pstats.Stats(pr)
    st.sort_stats('tottime')
    st.print_stats(14)
