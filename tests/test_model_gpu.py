"""BASELINE configs[2] at test size: the two hot paths inside the network they exist for.  One training step
(forward, loss, backward) of the reference's SimCLR ResNet-50 + MLP head (tests/carrier_resnet.py) with this library's
SyncBatchNorm and SupConLoss swapped in the way /root/reference/main_supcon.py:214,222-224 does it, against the same
network on torch.nn.BatchNorm2d and the reference's dense loss (oracle/dense_port.py); and the variant with ReLU /
add+ReLU folded into the batch-norm kernels."""
import copy

import pytest
import torch

from carrier_resnet import SupConResNet50, count_bn
from oracle.dense_port import dense_supcon_loss
from simclr_pytorch_distributed_b200 import SupConLoss, SyncBatchNorm

pytestmark = pytest.mark.gpu


def one_step(model, images, criterion):
    bsz = images.shape[0] // 2
    feats = model(images)                                          # main_supcon.py:266, normalised as :283
    f1, f2 = torch.split(feats, [bsz, bsz], dim=0)
    loss = criterion(torch.cat([f1.unsqueeze(1), f2.unsqueeze(1)], dim=1))     # :285-290
    model.zero_grad(set_to_none=True)
    loss.backward()
    return loss.detach().float(), {n: p.grad.detach().float().clone() for n, p in model.named_parameters()}


def rel(a, b):
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


@pytest.mark.parametrize('dtype,cl', [(torch.float32, False), (torch.bfloat16, True)])
def test_resnet50_step_with_both_drop_ins(dtype, cl):
    torch.manual_seed(0)
    dev = 'cuda'
    stock = SupConResNet50().to(dev)
    assert count_bn(stock) == 53
    if cl:
        stock = stock.to(memory_format=torch.channels_last)
    stock = stock.to(dtype)
    swapped = SyncBatchNorm.convert_sync_batchnorm(copy.deepcopy(stock))       # main_supcon.py:224
    fused = SupConResNet50(bn=SyncBatchNorm, fuse=True).to(dev)
    if cl:
        fused = fused.to(memory_format=torch.channels_last)
    fused = fused.to(dtype)
    fused.load_state_dict(stock.state_dict())
    assert sum(isinstance(m, SyncBatchNorm) for m in swapped.modules()) == 53
    images = torch.randn(64, 3, 32, 32, device=dev).to(dtype)
    if cl:
        images = images.contiguous(memory_format=torch.channels_last)
    ref_loss, ref_grads = one_step(stock, images, lambda f: dense_supcon_loss(f.float(), temperature=0.5).to(f.dtype))
    crit = SupConLoss(temperature=0.5)                             # main_supcon.py:214; engine 'auto'
    tol_loss, tol_grad = (1e-4, 2e-3) if dtype == torch.float32 else (2e-3 + 2 ** -8, 0.25)
    for name, model in (('swap', swapped), ('fused', fused)):
        loss, grads = one_step(model, images, crit)
        assert abs(float(loss) - float(ref_loss)) <= tol_loss * abs(float(ref_loss)), (name, float(loss), float(ref_loss))
        num = sum(float((grads[n] - ref_grads[n]).norm()) ** 2 for n in ref_grads) ** 0.5
        den = sum(float(ref_grads[n].norm()) ** 2 for n in ref_grads) ** 0.5
        assert num / den <= tol_grad, (name, num / den)
        # running statistics follow the stock module's
        sd, rd = dict(model.named_buffers()), dict(stock.named_buffers())
        worst = max(rel(sd[k].float(), rd[k].float()) for k in rd if k.endswith('running_var'))
        assert worst <= (1e-4 if dtype == torch.float32 else 2e-2), (name, worst)
    assert list(stock.state_dict().keys()) == list(swapped.state_dict().keys()) == list(fused.state_dict().keys())
