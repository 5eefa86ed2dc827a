import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from simclr_pytorch_distributed_b200 import SupConLoss
torch.manual_seed(8)
eager = SupConLoss(temperature=0.5, engine='bf16')
graphed = SupConLoss(temperature=0.5, engine='bf16', graph=True)
z = F.normalize(torch.randn(512, 2, 128, device='cuda'), dim=-1).to(torch.bfloat16)
def run(mod, scale, labels=None):
    a = z.clone().requires_grad_(True)
    l = mod(a, labels=labels)
    (scale * l).backward()
    return float(l), a.grad.float()
for labels in (None, torch.randint(0, 20, (512,), device='cuda')):
    l1, g1 = run(eager, 1.0, labels); l2, g2 = run(eager, 1.0, labels); l3, g3 = run(eager, 2.0, labels)
    l4, g4 = run(graphed, 1.0, labels); l5, g5 = run(graphed, 2.0, labels)
    print('labels' if labels is not None else 'simclr', 'eager twice maxdiff', float((g1 - g2).abs().max()), 'eager x2 vs 2*x1', float((g3 - 2 * g1).abs().max()),
          'graph vs eager', float((g4 - g1).abs().max()), 'graph x2 vs eager x2', float((g5 - g3).abs().max()), 'n diff', int((g5 != g3).sum()), 'max |g|', float(g1.abs().max()))
l4, g4 = run(graphed, 1.0); l4b, g4b = run(graphed, 1.0)
print('graph twice', float((g4 - g4b).abs().max()), int((g4 != g4b).sum()))
import os
print('PDL', os.environ.get('SCB_PDL'))
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    l6, g6 = run(eager, 1.0)
torch.cuda.synchronize()
l1, g1 = run(eager, 1.0)
print('eager other stream vs eager', float((g6 - g1).abs().max()), int((g6 != g1).sum()))
idx = (g4 != g1).nonzero()
print('where', idx[:8].tolist())
