"""Short driver for ncu: a few forward+backward passes of the loss at BASELINE config 4
(2N = 8192, D = 128, bf16 engine) through the C ABI.   python tools/prof_loss.py [iters] [B]"""
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from simclr_pytorch_distributed_b200 import _lib, losses as L   # noqa: E402


def main():
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 3
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
    engine = sys.argv[3] if len(sys.argv) > 3 else 'bf16'
    lib = _lib.load()
    dev = torch.device('cuda')
    V, D = 2, 128
    z = F.normalize(torch.randn(B, V, D, device=dev), dim=-1)
    z = z.to(torch.bfloat16) if engine == 'bf16' else z
    plan = L.LossPlan(B, V, D, engine=L._ENGINES[engine])
    k = torch.empty((plan.n_k, plan.ldk), dtype=plan.k_torch, device=dev)
    stats = torch.empty((4, plan.n_k), dtype=torch.float32, device=dev)
    loss = torch.empty(2, dtype=torch.float32, device=dev)
    g = torch.ones((), dtype=torch.float32, device=dev)
    dk = torch.empty((plan.rows_local, plan.ldk), dtype=torch.float32, device=dev)
    dz = torch.empty_like(z)
    prob = L._problem(plan, k, None, None, None, 0.5, 0.07)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    for it in range(iters):
        L.launch_prep(lib, z, plan, False, k, None, None)
        ev[0].record()
        L.launch_forward(lib, prob, plan, stats[0], stats[1], stats[2], loss, dev)
        ev[1].record()
        L.launch_backward(lib, prob, plan, stats[0], stats[2], g, dk, dev)
        ev[2].record()
        L.launch_finish(lib, dk, z, plan, False, None, dz)
        torch.cuda.synchronize()
    print('loss %.6f  fwd(+finalize) %.1f us  bwd(+init) %.1f us' % (
        loss[0].item(), ev[0].elapsed_time(ev[1]) * 1e3, ev[1].elapsed_time(ev[2]) * 1e3))


if __name__ == '__main__':
    main()
