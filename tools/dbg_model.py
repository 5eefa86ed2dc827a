import copy, sys, os
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests'))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from carrier_resnet import SupConResNet50
from oracle.dense_port import dense_supcon_loss
from simclr_pytorch_distributed_b200 import SupConLoss, SyncBatchNorm
from test_model_gpu import one_step, rel
torch.manual_seed(0)
for dtype, cl in ((torch.float32, False), (torch.float32, True), (torch.bfloat16, True)):
    stock = SupConResNet50().cuda()
    if cl: stock = stock.to(memory_format=torch.channels_last)
    stock = stock.to(dtype)
    swapped = SyncBatchNorm.convert_sync_batchnorm(copy.deepcopy(stock))
    stock2 = copy.deepcopy(stock)
    images = torch.randn(64, 3, 32, 32, device='cuda').to(dtype)
    if cl: images = images.contiguous(memory_format=torch.channels_last)
    dl = lambda f: dense_supcon_loss(f.float(), temperature=0.5).to(f.dtype)
    l0, g0 = one_step(stock, images, dl)
    l0b, g0b = one_step(stock2, images, dl)          # stock vs stock: run-to-run noise of the reference itself
    l1, g1 = one_step(swapped, images, dl)           # our BN, reference loss
    l2, g2 = one_step(copy.deepcopy(stock), images, SupConLoss(temperature=0.5))   # stock BN, our loss
    def tot(a, b):
        num = sum(float((a[n] - b[n]).norm()) ** 2 for n in b) ** 0.5
        den = sum(float(b[n].norm()) ** 2 for n in b) ** 0.5
        return num / den
    print(dtype, 'cl' if cl else 'nchw', 'loss', float(l0), float(l1), float(l2), 'grad rel: stock-vs-stock %.2e  ourBN %.2e  ourLoss %.2e' % (tot(g0b, g0), tot(g1, g0), tot(g2, g0)))
    worst = sorted(((rel(g1[n], g0[n]), n) for n in g0), reverse=True)[:4]
    print('   worst ourBN:', worst)
    first = [(n, '%.1e' % rel(g1[n], g0[n])) for n in list(g0)[-6:]]
    print('   last params:', first)
