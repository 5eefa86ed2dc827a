import copy, sys, os
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests'))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from carrier_resnet import SupConResNet50
from oracle.dense_port import dense_supcon_loss
from simclr_pytorch_distributed_b200 import SupConLoss, SyncBatchNorm
from test_model_gpu import one_step, rel
torch.manual_seed(0)
def tot(a, b):
    num = sum(float((a[n].double() - b[n].double()).norm()) ** 2 for n in b) ** 0.5
    den = sum(float(b[n].double().norm()) ** 2 for n in b) ** 0.5
    return num / den
for bsz in (64, 256):
    stock = SupConResNet50().cuda().to(memory_format=torch.channels_last)
    images = torch.randn(bsz, 3, 32, 32, device='cuda').contiguous(memory_format=torch.channels_last)
    dl = lambda f: dense_supcon_loss(f.float(), temperature=0.5).to(f.dtype)
    dl64 = lambda f: dense_supcon_loss(f, temperature=0.5)
    s64 = copy.deepcopy(stock).double()
    def step64(model, images):
        b = images.shape[0] // 2
        feats = model(images)
        f1, f2 = torch.split(feats, [b, b], dim=0)
        loss = dense_supcon_loss(torch.cat([f1.unsqueeze(1), f2.unsqueeze(1)], dim=1), temperature=0.5)
        model.zero_grad(set_to_none=True); loss.backward()
        return loss.detach(), {n: p.grad.detach().clone() for n, p in model.named_parameters()}
    l64, g64 = step64(s64, images.double())
    l0, g0 = one_step(copy.deepcopy(stock), images, dl)
    lp, gp = one_step(copy.deepcopy(stock), images * (1 + 1e-6), dl)
    swapped = SyncBatchNorm.convert_sync_batchnorm(copy.deepcopy(stock))
    l1, g1 = one_step(swapped, images, dl)
    print('bsz', bsz, 'loss fp64 %.6f stock32 %.6f ours32 %.6f' % (float(l64), float(l0), float(l1)))
    print('  grad rel vs fp64: stock32 %.2e  ours32 %.2e  | stock32 perturbed-input vs stock32 %.2e | ours32 vs stock32 %.2e' % (tot(g0, g64), tot(g1, g64), tot(gp, g0), tot(g1, g0)))
