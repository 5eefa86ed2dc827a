"""Quick timing of SyncBatchNorm fwd+bwd on BASELINE cfg 2 ([256,256,8,8] NHWC bf16): the two C-ABI launches in a CUDA
graph (no torch kernels in between) and through the module.   python tools/time_bn.py [N C H W]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from simclr_pytorch_distributed_b200 import SyncBatchNorm, _lib   # noqa: E402
from simclr_pytorch_distributed_b200.sync_batchnorm import _Kernels as K, _layout   # noqa: E402

shape = tuple(int(a) for a in sys.argv[1:5]) if len(sys.argv) >= 5 else (256, 256, 8, 8)
N, C, H, W = shape
dev = torch.device('cuda')
pool = 12
xs = [(torch.randn(shape, device=dev) * 1.5 + 0.5).to(torch.bfloat16).contiguous(memory_format=torch.channels_last) for _ in range(pool)]
dys = [torch.randn(shape, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last) for _ in range(pool)]
ys = [torch.empty_like(x) for x in xs]
dxs = [torch.empty_like(x) for x in xs]
w = torch.rand(C, device=dev) + 0.5
b = torch.randn(C, device=dev) * 0.1
rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
saved = torch.empty(2 * C + 1, device=dev)
out = torch.empty(4 * C, device=dev)
st = torch.empty(2 * C + 1, device=dev)
_, geom = _layout(xs[0])
fused = K.fused_supported(xs[0], geom, 1)


def raw(i, which):
    x, dy, y, dx = xs[i], dys[i], ys[i], dxs[i]
    if fused:
        if which & 1:
            assert K.forward_fused(x, y, geom, w, b, rm, rv, 0.1, 1e-5, saved, 0) == 0
        if which & 2:
            assert K.backward_fused(dy, x, dx, geom, saved, w, b, out[:2 * C], out[2 * C:3 * C], out[3 * C:], 0) == 0
    else:
        if which & 1:
            K.stats(x, geom, st)
            K.apply(x, y, geom, st, 1, w, b, rm, rv, 0.1, 1e-5, saved, 0)
        if which & 2:
            K.backward_reduce(dy, x, geom, saved, w, b, out[:2 * C], out[2 * C:3 * C], out[3 * C:], 0)
            K.backward_apply(dy, x, dx, geom, saved, w, b, out[:2 * C], 0)


def graph_time(fn, reps=20):
    for i in range(3):
        fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(pool):
            fn(i)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (reps * pool) * 1e3


t_f = graph_time(lambda i: raw(i, 1))
t_b = graph_time(lambda i: raw(i, 2))
t_fb = graph_time(lambda i: raw(i, 3))
bn = SyncBatchNorm(C).to(dev)
xg = [x.clone().requires_grad_(True) for x in xs]


def mod(i):
    xg[i].grad = None
    bn.weight.grad = None
    bn.bias.grad = None
    bn(xg[i]).backward(dys[i])


t_m = graph_time(mod)
S = N * C * H * W * 2
print('SCB_BN_MAX_CLUSTER=%s SCB_BN_FUSED=%s shape %s fused=%s: C-ABI graph fwd %.2f us, bwd %.2f us, fwd+bwd %.2f us (%.0f GB/s of 8S); module graph %.2f us (%.0f GB/s)' % (
    os.environ.get('SCB_BN_MAX_CLUSTER', '-'), os.environ.get('SCB_BN_FUSED', '1'), shape, fused, t_f, t_b, t_fb, 8 * S / t_fb / 1e3, t_m, 8 * S / t_m / 1e3))
