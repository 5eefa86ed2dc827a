#!/bin/bash
timeout 900 python -m pytest tests/test_supcon_gpu.py tests/test_supcon_parity_gpu.py tests/test_model_gpu.py -m gpu -q -x 2>&1 | tail -25 > gpurun_out/r2k_t.log
timeout 120 python - > gpurun_out/r2k_x3.log 2>&1 <<'PY'
import torch, torch.nn.functional as F, time
from simclr_pytorch_distributed_b200 import SupConLoss
z = torch.randn(4096, 2, 128, device='cuda')
pool = [z.clone().requires_grad_(True) for _ in range(8)]
crit = SupConLoss(temperature=0.5, engine='bf16x3', normalize=True)
one = torch.ones((), device='cuda')
def step(i):
    pool[i % 8].grad = None
    crit(pool[i % 8]).backward(one)
for i in range(3): step(i)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    for i in range(8): step(i)
g.replay(); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): g.replay()
e1.record(); torch.cuda.synchronize()
print('bf16x3 fp32 raw features + normalize: %.2f us/step' % (e0.elapsed_time(e1) / 160 * 1e3))
PY
