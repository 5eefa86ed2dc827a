"""SyncBatchNorm fwd+bwd over the layer shapes of the reference ResNet-50 (SURVEY.md Appendix B, N = 512 per GPU,
NHWC bf16) and BASELINE cfg 2, next to the ATen ops stock SyncBatchNorm runs (torch/nn/modules/_functions.py:39,
122,144,170) on the same tensors.  GB/s are the 8*S algorithmic bytes of SURVEY 8(d) over the CUDA-event time.

  python tools/bn_census.py [--iters 20] [--shapes all|small]
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

SHAPES = [(256, 256, 8, 8), (512, 512, 4, 4), (512, 256, 8, 8), (512, 2048, 4, 4), (512, 512, 8, 8), (512, 128, 16, 16),
          (512, 1024, 8, 8), (512, 256, 16, 16), (512, 64, 32, 32), (512, 512, 16, 16), (512, 128, 32, 32),
          (512, 256, 32, 32)]


def time_fn(fn, iters, n_sets):
    for i in range(3):
        fn(i % n_sets)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i % n_sets)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3     # microseconds


def census(iters=20, shapes=SHAPES, relu=False):
    from simclr_pytorch_distributed_b200 import SyncBatchNorm
    dev = torch.device('cuda')
    rows = []
    for shape in shapes:
        N, C, H, W = shape
        S = N * C * H * W * 2
        n_sets = max(2, min(16, int(400e6 // (4 * S)) + 1))      # x, dy, y, dx sets rotating through > L2 when they fit
        xs = [(torch.randn(shape, device=dev) * 1.5 + 0.5).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
              .requires_grad_(True) for _ in range(n_sets)]
        dys = [torch.randn(shape, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
               for _ in range(n_sets)]
        bn = SyncBatchNorm(C, fuse_relu=relu).to(dev)
        w, b = bn.weight.detach(), bn.bias.detach()

        def ours(i):
            x = xs[i]
            x.grad = None
            bn.weight.grad = None
            bn.bias.grad = None
            bn(x).backward(dys[i])

        count = torch.full((1,), N * H * W, dtype=torch.int32, device=dev)

        def aten(i):
            x, dy = xs[i].detach(), dys[i]
            mean, invstd = torch.batch_norm_stats(x, 1e-5)
            torch.batch_norm_elemt(x, w, b, mean, invstd, 1e-5)
            sum_dy, sum_dy_xmu, _, _ = torch.batch_norm_backward_reduce(dy, x, mean, invstd, w, True, True, True)
            torch.batch_norm_backward_elemt(dy, x, mean, invstd, w, sum_dy, sum_dy_xmu, count)

        t_ours = time_fn(ours, iters, n_sets)
        t_aten = time_fn(aten, iters, n_sets)
        # the same step as a CUDA graph (what a captured training step pays: no host launch gaps)
        g = torch.cuda.CUDAGraph()
        try:
            with torch.cuda.graph(g):
                for i in range(n_sets):
                    ours(i)
            t_graph = time_fn(lambda i: g.replay(), max(2, iters // n_sets), 1) / n_sets
        except Exception as e:      # say so
            t_graph = None
            print('graph capture failed for', shape, type(e).__name__, e, file=sys.stderr)
            torch.cuda.synchronize()
        rows.append({'shape': list(shape), 'MB': S / 1e6, 'ours_us': t_ours, 'ours_graph_us': t_graph, 'aten_us': t_aten,
                     'ours_GBs': 8 * S / (t_graph or t_ours) / 1e3, 'aten_GBs': 8 * S / t_aten / 1e3})
        del xs, dys
        torch.cuda.empty_cache()
    return rows


if __name__ == '__main__':
    ap = argparse.ArgumentParser()
    ap.add_argument('--iters', type=int, default=20)
    ap.add_argument('--shapes', default='all')
    ap.add_argument('--relu', action='store_true')
    a = ap.parse_args()
    shapes = SHAPES if a.shapes == 'all' else SHAPES[:3]
    for r in census(a.iters, shapes, a.relu):
        print(json.dumps(r), flush=True)
