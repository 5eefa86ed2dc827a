import cProfile, pstats, torch, torch.nn.functional as F, time, sys
sys.path.insert(0, '/root/repo')
from simclr_pytorch_distributed_b200 import SupConLoss
z = [F.normalize(torch.randn(4096, 2, 128, device='cuda'), dim=-1).to(torch.bfloat16).requires_grad_(True) for _ in range(4)]
crit = SupConLoss(temperature=0.5)
def step(i):
    f = z[i % 4]; f.grad = None
    crit(f).backward()
for i in range(20): step(i)
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(300): step(i)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print('host time per eager step %.1f us (+%.1f us drain)' % ((t1 - t0) / 300 * 1e6, (t2 - t1) * 1e6))
pr = cProfile.Profile(); pr.enable()
for i in range(300): step(i)
pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats('tottime').print_stats(18)
