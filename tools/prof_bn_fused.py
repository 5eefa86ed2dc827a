"""ncu driver: fused SyncBN forward/backward on cfg 2 through the module API.  python tools/prof_bn_fused.py [iters]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from simclr_pytorch_distributed_b200 import SyncBatchNorm   # noqa: E402

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 10
dev = torch.device('cuda')
bn = SyncBatchNorm(256).to(dev)
xs = [(torch.randn(256, 256, 8, 8, device=dev) * 1.5 + 0.5).to(torch.bfloat16).contiguous(
    memory_format=torch.channels_last).requires_grad_(True) for _ in range(12)]
dys = [torch.randn(256, 256, 8, 8, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
       for _ in range(12)]
for it in range(iters):
    x = xs[it % 12]
    x.grad = None
    bn(x).backward(dys[it % 12])
torch.cuda.synchronize()
print('ok')
