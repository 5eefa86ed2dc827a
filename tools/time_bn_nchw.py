"""SyncBatchNorm fwd+bwd on NCHW tensors (what the reference model runs without channels_last), ours vs the ATen ops
stock SyncBatchNorm launches.   python tools/time_bn_nchw.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from simclr_pytorch_distributed_b200 import SyncBatchNorm   # noqa: E402

dev = torch.device('cuda')
SHAPES = [(256, 256, 8, 8), (512, 64, 32, 32), (512, 256, 32, 32), (512, 512, 16, 16), (512, 2048, 4, 4)]


def time_fn(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


for dtype in (torch.float32, torch.bfloat16):
    for shape in SHAPES:
        N, C, H, W = shape
        x = torch.randn(shape, device=dev).to(dtype).requires_grad_(True)
        dy = torch.randn(shape, device=dev).to(dtype)
        bn = SyncBatchNorm(C).to(dev)
        w, b = bn.weight.detach(), bn.bias.detach()
        count = torch.full((1,), N * H * W, dtype=torch.int32, device=dev)

        def ours():
            x.grad = None
            bn.weight.grad = None
            bn.bias.grad = None
            bn(x).backward(dy)

        def aten():
            xd = x.detach()
            mean, invstd = torch.batch_norm_stats(xd, 1e-5)
            torch.batch_norm_elemt(xd, w, b, mean, invstd, 1e-5)
            sd, sx, _, _ = torch.batch_norm_backward_reduce(dy, xd, mean, invstd, w, True, True, True)
            torch.batch_norm_backward_elemt(dy, xd, mean, invstd, w, sd, sx, count)

        S = x.numel() * x.element_size()
        to, ta = time_fn(ours), time_fn(aten)
        print('%s NCHW %s: ours %.1f us (%.0f GB/s of 8S), ATen ops %.1f us (%.0f GB/s)  ratio %.2f' % (
            str(dtype).split('.')[-1], shape, to, 8 * S / to / 1e3, ta, 8 * S / ta / 1e3, ta / to))
