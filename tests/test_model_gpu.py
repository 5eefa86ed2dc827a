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


def total_rel(a, b):
    num = sum(float((a[n].double() - b[n].double()).norm()) ** 2 for n in b) ** 0.5
    den = sum(float(b[n].double().norm()) ** 2 for n in b) ** 0.5
    return num / den


def build(dtype, cl, **kw):
    torch.manual_seed(0)
    stock = SupConResNet50(**kw).to('cuda')
    if cl:
        stock = stock.to(memory_format=torch.channels_last)
    stock = stock.to(dtype)
    swapped = SyncBatchNorm.convert_sync_batchnorm(copy.deepcopy(stock))       # main_supcon.py:224
    fused = SupConResNet50(bn=SyncBatchNorm, fuse=True, **kw).to('cuda')
    if cl:
        fused = fused.to(memory_format=torch.channels_last)
    fused = fused.to(dtype)
    fused.load_state_dict(stock.state_dict())
    return stock, swapped, fused


def dense(f):
    return dense_supcon_loss(f.float(), temperature=0.5).to(f.dtype)


@pytest.mark.parametrize('dtype,cl', [(torch.float32, False), (torch.float32, True), (torch.bfloat16, True)])
def test_small_carrier_step_gradients(dtype, cl):
    """A shallow, narrow version of the carrier (one Bottleneck per stage, width 16): shallow enough for the
    parameter gradients of two arithmetically different batch-norm implementations to be comparable."""
    stock, swapped, fused = build(dtype, cl, width=16, blocks=(1, 1, 1, 1))
    images = torch.randn(128, 3, 32, 32, device='cuda').to(dtype)
    if cl:
        images = images.contiguous(memory_format=torch.channels_last)
    truth = copy.deepcopy(stock).double()
    t_loss, t_grads = one_step(truth, images.double(), lambda f: dense_supcon_loss(f, temperature=0.5))
    s_loss, s_grads = one_step(stock, images, dense)
    noise = total_rel(s_grads, t_grads)                            # what the stock modules themselves lose vs fp64
    crit = SupConLoss(temperature=0.5)                             # main_supcon.py:214; engine 'auto'
    for name, model in (('swap', swapped), ('fused', fused)):
        loss, grads = one_step(model, images, crit)
        tol_loss = 1e-4 if dtype == torch.float32 else 2e-3 + 2 ** -7
        assert abs(float(loss) - float(s_loss)) <= tol_loss * abs(float(s_loss)), (name, float(loss), float(s_loss))
        err = total_rel(grads, t_grads)
        assert err <= 1.5 * noise + (1e-4 if dtype == torch.float32 else 1e-2), (name, err, noise)


@pytest.mark.parametrize('dtype,cl', [(torch.float32, False), (torch.bfloat16, True)])
def test_resnet50_step_with_both_drop_ins(dtype, cl):
    """The full 53-BN network.  At random initialisation its parameter gradients are chaotic -- the stock fp32 modules
    are ~0.5 (relative, Frobenius) away from their own fp64 evaluation, and a 1e-6 perturbation of the input moves
    them by ~0.4 -- so the gradients are held to "no further from the fp64 truth than the stock modules are", and the
    loss, running statistics and state-dict layout to the stock modules directly."""
    stock, swapped, fused = build(dtype, cl)
    assert count_bn(stock) == 53
    assert sum(isinstance(m, SyncBatchNorm) for m in swapped.modules()) == 53
    images = torch.randn(64, 3, 32, 32, device='cuda').to(dtype)
    if cl:
        images = images.contiguous(memory_format=torch.channels_last)
    truth = copy.deepcopy(stock).double()
    t_loss, t_grads = one_step(truth, images.double(), lambda f: dense_supcon_loss(f, temperature=0.5))
    ref_loss, ref_grads = one_step(stock, images, dense)
    noise = total_rel(ref_grads, t_grads)
    crit = SupConLoss(temperature=0.5)
    tol_loss = 5e-4 if dtype == torch.float32 else 2e-2
    for name, model in (('swap', swapped), ('fused', fused)):
        loss, grads = one_step(model, images, crit)
        assert abs(float(loss) - float(t_loss)) <= tol_loss * abs(float(t_loss)), (name, float(loss), float(t_loss))
        err = total_rel(grads, t_grads)
        assert err <= 1.3 * noise + 0.05, (name, err, noise)
        sd, rd = dict(model.named_buffers()), dict(stock.named_buffers())
        worst = max(rel(sd[k].float(), rd[k].float()) for k in rd if k.endswith('running_var'))
        assert worst <= (1e-3 if dtype == torch.float32 else 5e-2), (name, worst)
    assert list(stock.state_dict().keys()) == list(swapped.state_dict().keys()) == list(fused.state_dict().keys())


def test_small_carrier_with_projection_tail():
    """The fused arm of `bench.py --workload model`: ReLU / add+ReLU inside the batch-norm kernels and the end of the
    head (Linear, normalise, view stacking) through ProjectionTail, whose bf16 output the loss uses in place."""
    dtype, kw = torch.bfloat16, dict(width=16, blocks=(1, 1, 1, 1))
    stock, _, _ = build(dtype, True, **kw)
    tailed = SupConResNet50(bn=SyncBatchNorm, fuse=True, tail=True, **kw).to('cuda').to(
        memory_format=torch.channels_last).to(dtype)
    tailed.load_state_dict(stock.state_dict())
    images = torch.randn(128, 3, 32, 32, device='cuda').to(dtype).contiguous(memory_format=torch.channels_last)
    truth = copy.deepcopy(stock).double()
    t_loss, t_grads = one_step(truth, images.double(), lambda f: dense_supcon_loss(f, temperature=0.5))
    s_loss, s_grads = one_step(stock, images, dense)
    noise = total_rel(s_grads, t_grads)
    crit = SupConLoss(temperature=0.5)
    z = tailed(images)
    assert z.shape == (64, 2, 128) and z.is_contiguous()
    loss = crit(z)
    tailed.zero_grad(set_to_none=True)
    loss.backward()
    grads = {n: p.grad.detach().float().clone() for n, p in tailed.named_parameters()}
    assert abs(float(loss) - float(s_loss)) <= (2e-3 + 2 ** -7) * abs(float(s_loss))
    err = total_rel(grads, t_grads)
    assert err <= 1.5 * noise + 1e-2, (err, noise)
    assert list(tailed.state_dict().keys()) == list(stock.state_dict().keys())
