import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from simclr_pytorch_distributed_b200 import SupConLoss
torch.manual_seed(8)
eager = SupConLoss(temperature=0.5, engine='bf16')
graphed = SupConLoss(temperature=0.5, engine='bf16', graph=True)
bufs = [torch.empty(512, 2, 128, dtype=torch.bfloat16, device='cuda') for _ in range(2)]
for step in range(7):
    z = F.normalize(torch.randn(512, 2, 128, device='cuda'), dim=-1).to(torch.bfloat16)
    if step < 4:
        bufs[step % 2].copy_(z); src = bufs[step % 2]
    else:
        src = z.clone()
    labels = torch.randint(0, 20, (512,), device='cuda') if step % 3 == 2 else None
    a = src.detach().clone().requires_grad_(True)
    b = src.detach().requires_grad_(True)
    la = eager(a, labels=labels); (2.0 * la).backward()
    lb = graphed(b, labels=labels); (2.0 * lb).backward()
    print(step, 'labels' if labels is not None else 'simclr', float(la), float(lb), float((a.grad.float() - b.grad.float()).abs().max()), float(a.grad.float().abs().max()), float(b.grad.float().abs().max()))
