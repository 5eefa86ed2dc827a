"""Quick timing of the loss at BASELINE config 4 (2N = 8192, D = 128, bf16 features in place): CUDA-graph replay of
fwd+bwd through the module API, and the two C-ABI calls timed apart.   python tools/time_loss.py [B] [reps]"""
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from simclr_pytorch_distributed_b200 import SupConLoss, _lib   # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
lib = _lib.load()
pool = 16
zs = [F.normalize(torch.randn(B, 2, 128, device='cuda'), dim=-1).to(torch.bfloat16).requires_grad_(True) for _ in range(pool)]
crit = SupConLoss(temperature=0.5, engine='bf16')
one = torch.ones((), dtype=torch.bfloat16, device='cuda')


def step(i):
    z = zs[i % pool]
    z.grad = None
    crit(z).backward(one)


for i in range(3):
    step(i)
torch.cuda.synchronize()
c0 = lib.scb_launch_count()
step(0)
launches = lib.scb_launch_count() - c0
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    for i in range(pool):
        step(i)
g.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    g.replay()
e1.record()
torch.cuda.synchronize()
per = e0.elapsed_time(e1) / (reps * pool) * 1e3
# forward and backward apart (eager, host running ahead)
ef = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
tf = tb = 0.0
for it in range(10):
    z = zs[it % pool]
    z.grad = None
    torch.cuda._sleep(300000)
    ef[0].record()
    l = crit(z)
    ef[1].record()
    l.backward(one)
    ef[2].record()
    torch.cuda.synchronize()
    if it >= 2:
        tf += ef[0].elapsed_time(ef[1]); tb += ef[1].elapsed_time(ef[2])
# eager steps back to back (host-bound: what a training loop without graphs pays per step)
import time
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(100):
    step(i)
torch.cuda.synchronize()
eager_us = (time.perf_counter() - t0) / 100 * 1e6
print('SCB_PDL=%s SCB_BWD_DIRECT=%s B=%d launches=%d graph step %.2f us  eager step %.1f us (apart: fwd %.1f us, bwd %.1f us)  loss %.4f' % (
    os.environ.get('SCB_PDL', '1'), os.environ.get('SCB_BWD_DIRECT', '1'), B, launches, per, eager_us, tf / 8 * 1e3, tb / 8 * 1e3,
    float(crit(zs[0].detach()))))
