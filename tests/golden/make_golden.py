"""Generate the golden input/output vectors under tests/golden/ by running the UNMODIFIED
reference (`/root/reference/losses.py`) -- and, for SyncBN, `torch.nn.BatchNorm2d` on the
rank-concatenated batch, because the arithmetic of `nn.SyncBatchNorm` lives in CUDA-only ATen
ops that cannot run in a GPU-less container.

Run in the build container only (needs /root/reference):   python tests/golden/make_golden.py
The .npz files it writes are committed; nothing at test/bench time reads /root/reference.
"""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, '/root/reference')
from losses import SupConLoss  # noqa: E402  (the reference, imported as is)

torch.manual_seed(20261106)
torch.set_num_threads(4)


def unit(*shape, dtype=torch.float64):
    z = torch.randn(*shape, dtype=dtype)
    # a shared direction gives the similarity matrix some structure
    z = z + 0.5 * torch.randn(1, 1, shape[-1], dtype=dtype)
    return F.normalize(z, dim=-1)


def run_loss(name, feats, labels=None, mask=None, T=0.07, mode='all', Tb=0.07):
    f = feats.clone().requires_grad_(True)
    crit = SupConLoss(temperature=T, contrast_mode=mode, base_temperature=Tb)
    loss = crit(f, labels=labels, mask=mask)
    loss.backward()
    out = dict(features=feats.numpy(), loss=loss.detach().numpy(), grad=f.grad.numpy(),
               temperature=np.float64(T), base_temperature=np.float64(Tb), mode=np.array(mode))
    if labels is not None:
        out['labels'] = labels.numpy()
    if mask is not None:
        out['mask'] = mask.numpy()
    np.savez_compressed(os.path.join(HERE, f'loss_{name}.npz'), **out)
    print(f'loss_{name}: loss={float(loss):.10f}')


def loss_cases():
    run_loss('simclr_all_f64', unit(32, 2, 16), T=0.5)
    run_loss('simclr_all_t007', unit(20, 2, 24), T=0.07)
    run_loss('simclr_one', unit(16, 2, 16), T=0.5, mode='one')
    lab = torch.randint(0, 5, (24,))
    run_loss('labels_all', unit(24, 2, 8), labels=lab, T=0.07)
    run_loss('labels_one', unit(24, 2, 8), labels=lab, T=0.1, mode='one', Tb=0.2)
    m = (torch.rand(12, 12, dtype=torch.float64) < 0.4).double() * torch.rand(12, 12, dtype=torch.float64)
    m = m + torch.eye(12, dtype=torch.float64)          # every anchor keeps >= 1 positive
    run_loss('mask_asym_all_v3', unit(12, 3, 8), mask=m, T=0.2)
    run_loss('mask_asym_one_v3', unit(12, 3, 8), mask=m, T=0.2, mode='one')
    lab1 = torch.arange(10).repeat(2)                    # V=1: positives only through labels
    run_loss('labels_v1', unit(20, 1, 8), labels=lab1, T=0.3)
    run_loss('highdim', unit(8, 2, 2, 4).reshape(8, 2, 2, 4), T=0.5)
    # un-normalised features: the loss must not normalise by itself
    run_loss('unnormalised', 0.7 * torch.randn(16, 2, 8, dtype=torch.float64), T=0.5)
    # BASELINE.json configs[0]: [256, 2, 128], temp 0.5, fp32 (the reference's CPU-runnable case)
    g = torch.Generator().manual_seed(1234)
    z = F.normalize(torch.randn(256, 2, 128, generator=g), dim=-1)
    run_loss('cfg1_f32', z, T=0.5)
    # BASELINE.json configs[4] in miniature: labels, 100 classes, all vs one
    lab = torch.randint(0, 100, (256,), generator=g)
    z = F.normalize(torch.randn(256, 2, 128, generator=g), dim=-1)
    run_loss('cfg5mini_all_f32', z, labels=lab, T=0.07)
    run_loss('cfg5mini_one_f32', z, labels=lab, T=0.07, mode='one')


def glue_case(W, bsz, D, T):
    """main_supcon.py:266-290 for W ranks (the reference hard-wires W=2; for W=2 the ordering below
    is exactly its [fea00, fea10, fea01, fea11])."""
    parts = [torch.randn(2 * bsz, D, dtype=torch.float64).requires_grad_(True) for _ in range(W)]
    v0 = [p[:bsz] for p in parts]
    v1 = [p[bsz:] for p in parts]
    feats = torch.cat(v0 + v1, dim=0)
    n_fea = F.normalize(feats, dim=1)
    f1, f2 = torch.split(n_fea, [W * bsz, W * bsz], dim=0)
    n_features = torch.cat([f1.unsqueeze(1), f2.unsqueeze(1)], dim=1)
    loss = SupConLoss(temperature=T)(n_features)
    loss.backward()
    out = {f'part{r}': parts[r].detach().numpy() for r in range(W)}
    out.update({f'grad{r}': parts[r].grad.numpy() for r in range(W)})
    out.update(loss=loss.detach().numpy(), n_features=n_features.detach().numpy(),
               temperature=np.float64(T), world=np.int64(W), bsz=np.int64(bsz))
    np.savez_compressed(os.path.join(HERE, f'glue_w{W}.npz'), **out)
    print(f'glue_w{W}: loss={float(loss):.10f}')


def bn_case(name, counts, C, hw, dtype=torch.float64, momentum=0.1, eps=1e-5):
    xs = [(torch.randn(n, C, *hw, dtype=dtype) * (0.5 + 1.5 * torch.rand(1, C, 1, 1, dtype=dtype))
           + 4.0 * (torch.rand(1, C, 1, 1, dtype=dtype) - 0.5)) for n in counts]
    dys = [torch.randn_like(x) for x in xs]
    bn = torch.nn.BatchNorm2d(C, eps=eps, momentum=momentum).to(dtype)
    with torch.no_grad():
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.normal_(0, 0.1)
        bn.running_mean.normal_(0, 0.3)
        bn.running_var.uniform_(0.5, 2.0)
    rm0, rv0 = bn.running_mean.clone(), bn.running_var.clone()
    x = torch.cat(xs, 0).requires_grad_(True)
    y = bn(x)
    y.backward(torch.cat(dys, 0))
    out = dict(weight=bn.weight.detach().numpy(), bias=bn.bias.detach().numpy(),
               running_mean0=rm0.numpy(), running_var0=rv0.numpy(),
               running_mean1=bn.running_mean.numpy(), running_var1=bn.running_var.numpy(),
               momentum=np.float64(momentum), eps=np.float64(eps), counts=np.array(counts),
               grad_weight=bn.weight.grad.numpy(), grad_bias=bn.bias.grad.numpy())
    ys = torch.split(y.detach(), counts, 0)
    dxs = torch.split(x.grad, counts, 0)
    for r in range(len(counts)):
        out[f'x{r}'] = xs[r].numpy()
        out[f'dy{r}'] = dys[r].numpy()
        out[f'y{r}'] = ys[r].numpy()
        out[f'dx{r}'] = dxs[r].numpy()
    np.savez_compressed(os.path.join(HERE, f'bn_{name}.npz'), **out)
    print(f'bn_{name}: ok')


if __name__ == '__main__':
    loss_cases()
    glue_case(2, 6, 16, 0.5)
    glue_case(4, 4, 16, 0.5)
    bn_case('w1', [8], 6, (3, 3))
    bn_case('w2', [4, 4], 6, (3, 3))
    bn_case('w3_ragged', [3, 5, 2], 5, (2, 4))
    bn_case('w2_f32', [16, 16], 32, (4, 4), dtype=torch.float32)
