"""Short driver for ncu: the four SyncBatchNorm kernels on BASELINE config 2 ([256,256,8,8] NHWC
bf16) through the C ABI, rotating over more buffers than fit in L2.   python tools/prof_bn.py [iters]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from simclr_pytorch_distributed_b200.sync_batchnorm import _Kernels as K, _layout   # noqa: E402


def main():
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    shape = tuple(int(v) for v in sys.argv[2].split(',')) if len(sys.argv) > 2 else (256, 256, 8, 8)
    N, C, H, W = shape
    dev = torch.device('cuda')
    pool = max(2, min(12, int(600e6 // (N * C * H * W * 2 * 4))))
    xs = [(torch.randn(N, C, H, W, device=dev) * 1.5 + 0.5).to(torch.bfloat16).contiguous(
        memory_format=torch.channels_last) for _ in range(pool)]
    dys = [torch.randn(N, C, H, W, device=dev).to(torch.bfloat16).contiguous(
        memory_format=torch.channels_last) for _ in range(pool)]
    ys = [torch.empty_like(x) for x in xs]
    dxs = [torch.empty_like(x) for x in xs]
    w = torch.rand(C, device=dev) + 0.5
    b = torch.randn(C, device=dev) * 0.1
    rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    st = torch.empty(2 * C + 1, device=dev)
    saved = torch.empty(2 * C + 1, device=dev)
    out = torch.empty(4 * C, device=dev)
    _, geom = _layout(xs[0])
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
    tot = [0.0] * 4
    for it in range(iters):
        x, dy, y, dx = xs[it % pool], dys[it % pool], ys[it % pool], dxs[it % pool]
        torch.cuda._sleep(400000)
        ev[0].record()
        K.stats(x, geom, st)
        ev[1].record()
        K.apply(x, y, geom, st, 1, w, b, rm, rv, 0.1, 1e-5, saved, 0)
        ev[2].record()
        K.backward_reduce(dy, x, geom, saved, w, b, out[:2 * C], out[2 * C:3 * C], out[3 * C:], 0)
        ev[3].record()
        K.backward_apply(dy, x, dx, geom, saved, w, b, out[:2 * C], 0)
        ev[4].record()
        torch.cuda.synchronize()
        if it >= 3:
            for j in range(4):
                tot[j] += ev[j].elapsed_time(ev[j + 1])
    n = max(1, iters - 3)
    print('shape', shape, 'us per kernel: stats %.2f apply %.2f bwd_reduce %.2f bwd_apply %.2f' % tuple(t / n * 1e3 for t in tot))


if __name__ == '__main__':
    main()
