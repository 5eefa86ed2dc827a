"""torch port of the reference's *dense* algorithm, used only as the timed baseline (`bench.py` cpu_baseline /
`--impl reference` on the host cores, `gpu_baseline` / `parity` in eager PyTorch on the GPU: the on-box bar of
SURVEY.md F12) and as a second checker in tests.
TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Why a port: the reference is pure Python on PyTorch; /root/reference does not
exist on the GPU box, so the reference's CPU path is restated here op for op -- the same dense
[A x C] temporaries and autograd backward as /root/reference/losses.py:52-91 -- so that its cost
on the box's host cores is what the reference would cost there.  The SyncBN baseline is
`torch.nn.functional.batch_norm` in training mode, which is what `nn.SyncBatchNorm` itself
runs on a CPU / single rank (torch/nn/modules/batchnorm.py:818-828).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def dense_supcon_loss(features, labels=None, mask=None, temperature=0.07,
                      contrast_mode='all', base_temperature=0.07):
    """Dense SupCon/NT-Xent loss with the reference's temporaries (losses.py:17-93)."""
    if features.dim() < 3:
        raise ValueError('`features` needs to be [bsz, n_views, ...],'
                         'at least 3 dimensions are required')
    feats = features.reshape(features.shape[0], features.shape[1], -1)      # :36-37
    n_samples, n_views = feats.shape[0], feats.shape[1]
    if labels is not None and mask is not None:                             # :40-41
        raise ValueError('Cannot define both `labels` and `mask`')
    if labels is None and mask is None:                                     # :42-43
        base = torch.eye(n_samples, dtype=torch.float32, device=feats.device)
    elif labels is not None:                                                # :44-48
        col = labels.contiguous().view(-1, 1)
        if col.shape[0] != n_samples:
            raise ValueError('Num of labels does not match num of features')
        base = (col == col.T).float()
    else:                                                                   # :49-50
        base = mask.float()
    contrast = torch.cat(torch.unbind(feats, dim=1), dim=0)                 # :53
    if contrast_mode == 'one':                                              # :54-56
        anchors, n_anchor_views = feats[:, 0], 1
    elif contrast_mode == 'all':                                            # :57-59
        anchors, n_anchor_views = contrast, n_views
    else:
        raise ValueError('Unknown mode: {}'.format(contrast_mode))
    sim = (anchors @ contrast.T) / temperature                              # :64-66
    sim = sim - sim.max(dim=1, keepdim=True).values.detach()                # :68-69
    tiled = base.repeat(n_anchor_views, n_views)                            # :72
    keep = torch.ones_like(tiled)                                           # :74-79
    rows = torch.arange(n_samples * n_anchor_views, device=feats.device)
    keep[rows, rows] = 0
    tiled = tiled * keep                                                    # :80
    expo = torch.exp(sim) * keep                                            # :83
    logp = sim - torch.log(expo.sum(dim=1, keepdim=True))                   # :84
    pos_mean = (tiled * logp).sum(dim=1) / tiled.sum(dim=1)                 # :87
    per_anchor = -(temperature / base_temperature) * pos_mean               # :90
    return per_anchor.view(n_anchor_views, n_samples).mean()                # :91


def dense_supcon_fwd_bwd(features, labels=None, mask=None, **kw):
    """One forward + backward w.r.t. `features`; returns (loss, grad)."""
    f = features.detach().clone().requires_grad_(True)
    loss = dense_supcon_loss(f, labels, mask, **kw)
    loss.backward()
    return loss.detach(), f.grad


def bn_train_fwd_bwd(x, dy, weight, bias, running_mean, running_var, momentum=0.1, eps=1e-5):
    """Training-mode batch norm forward+backward on one (or the rank-concatenated) batch:
    the CPU path of `nn.SyncBatchNorm` (torch/nn/modules/batchnorm.py:818-828)."""
    x = x.detach().clone().requires_grad_(True)
    w = weight.detach().clone().requires_grad_(True)
    b = bias.detach().clone().requires_grad_(True)
    y = F.batch_norm(x, running_mean, running_var, w, b, True, momentum, eps)
    y.backward(dy)
    return y.detach(), x.grad, w.grad, b.grad
