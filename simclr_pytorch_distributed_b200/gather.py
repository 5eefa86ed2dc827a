"""Row-sharded global-batch loss: the W-rank replacement of the reference's gather / normalise
glue plus criterion call, /root/reference/main_supcon.py:266-290.

The reference all-gathers the embeddings and then lets EVERY rank compute the full 2N x 2N loss
(and is hard-wired to 2 ranks, main_supcon.py:271-279).  Here each rank

  1. normalises its own rows (fused `F.normalize`, main_supcon.py:283) and all-gathers them (NCCL),
  2. runs the forward for ITS row block against all columns,
  3. all-gathers the per-row (log-sum-exp, positive count) and its share of the loss,

and returns the same global loss on every rank.  The backward needs no collective: with the
symmetric form of the gradient a rank computes exactly
d L_global / d (its own embeddings) -- what the reference obtains by re-inserting its local slice
into the gathered list (main_supcon.py:271-274) -- so DDP's parameter averaging works unchanged.

Row order of the gathered matrix: rank-major, view-major inside a rank (include/simclr_b200.h).
The loss is invariant under a consistent re-ordering of the samples, so this matches the reference's
[view 0 of every rank ; view 1 of every rank] ordering (main_supcon.py:276-279).
"""
from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn as nn

from . import _lib
from . import losses as L


class _CudaOps:
    """The four launches of the single-GPU path, re-used on the sharded problem."""
    requires_cuda = True
    prep = staticmethod(L.launch_prep)
    forward = staticmethod(L.launch_forward)
    backward = staticmethod(L.launch_backward)
    finish = staticmethod(L.launch_finish)
    backward_features = staticmethod(L.launch_backward_features)

    @staticmethod
    def lib():
        return _lib.load()


OPS = _CudaOps      # tests replace this with an oracle-backed stand-in to exercise the host logic on CPU


def _all_gather_flat(out, inp, group):
    """all_gather_into_tensor, or the list form on backends without it (gloo)."""
    if dist.get_backend(group) != 'gloo':
        dist.all_gather_into_tensor(out, inp, group=group)
    else:
        world = dist.get_world_size(group)
        parts = list(out.view(world, -1).unbind(0))
        dist.all_gather(parts, inp.reshape(-1), group=group)


class _DistSupConFunction(torch.autograd.Function):

    @staticmethod
    def forward(ctx, feats, labels_all, plan, temperature, base_temperature, normalize, group):
        ops = OPS
        lib = ops.lib()
        dev = feats.device
        rows, n_k, ldk = plan.rows_local, plan.n_k, plan.ldk
        world = plan.n_seg
        k_local = torch.empty((rows, ldk), dtype=plan.k_torch, device=dev)
        inv_norm = torch.empty(rows, dtype=torch.float32, device=dev) if normalize else None
        ops.prep(lib, feats, plan, normalize, k_local, None, inv_norm)
        k_all = torch.empty((n_k, ldk), dtype=plan.k_torch, device=dev)
        _all_gather_flat(k_all, k_local, group)                                  # exchange 1: embeddings
        # local row statistics: [lse(rows) | npos(rows) | loss share, 3 pad]
        packed = torch.empty(2 * rows + 4, dtype=torch.float32, device=dev)
        possum = torch.empty(rows, dtype=torch.float32, device=dev)
        prob = L._problem(plan, k_all, None, labels_all, None, temperature, base_temperature)
        ops.forward(lib, prob, plan, packed[0:rows], possum, packed[rows:2 * rows], packed[2 * rows:2 * rows + 1], dev)
        gathered = torch.empty((world, 2 * rows + 4), dtype=torch.float32, device=dev)
        _all_gather_flat(gathered, packed, group)                               # exchange 2: row statistics
        lse_all = gathered[:, 0:rows].contiguous().view(-1)
        npos_all = gathered[:, rows:2 * rows].contiguous().view(-1)
        loss = gathered[:, 2 * rows].sum()           # every rank's share of mean_i loss_i (losses.py:91)
        ctx.plan, ctx.args = plan, (temperature, base_temperature, normalize)
        ctx.save_for_backward(feats, labels_all, k_all, lse_all, npos_all, inv_norm)
        return loss.to(feats.dtype) if feats.dtype != torch.float32 else loss

    @staticmethod
    def backward(ctx, grad_out):
        ops = OPS
        lib = ops.lib()
        feats, labels_all, k_all, lse_all, npos_all, inv_norm = ctx.saved_tensors
        plan = ctx.plan
        temperature, base_temperature, normalize = ctx.args
        dev = feats.device
        g = L.grad_scalar(grad_out, feats.dtype)
        prob = L._problem(plan, k_all, None, labels_all, None, temperature, base_temperature)
        dfeats = torch.empty(feats.shape, dtype=feats.dtype, device=dev)
        if plan.engine != _lib.ENGINE_FP32 and hasattr(ops, 'backward_features'):
            ops.backward_features(lib, prob, plan, lse_all, npos_all, g, feats, normalize, inv_norm, dfeats, dev)
        else:                                                                    # no collective needed either way
            dk = torch.empty((plan.rows_local, plan.ldk), dtype=torch.float32, device=dev)
            ops.backward(lib, prob, plan, lse_all, npos_all, g, dk, dev)
            ops.finish(lib, dk, feats, plan, normalize, inv_norm, dfeats)
        return dfeats, None, None, None, None, None, None


class DistributedSupConLoss(nn.Module):
    """SupCon / NT-Xent over the global batch with the rows sharded across the ranks.

    forward(features, labels=None): `features` [bsz, n_views, ...] are THIS rank's embeddings (raw
    model outputs; `normalize=True` applies main_supcon.py:283), `labels` [bsz] this rank's labels.
    Every rank must pass the same bsz and n_views.  Returns the global loss (identical on all ranks);
    its gradient with respect to `features` is d L_global / d features, as in the reference.
    Explicit `mask` arguments are a single-GPU feature of `SupConLoss` (the reference cannot shard
    them either).
    """

    def __init__(self, temperature=0.07, contrast_mode='all', base_temperature=0.07, engine='auto',
                 normalize=True, process_group=None):
        super().__init__()
        self.temperature = temperature
        self.contrast_mode = contrast_mode
        self.base_temperature = base_temperature
        self.engine = engine
        self.normalize = normalize
        self.process_group = process_group

    def forward(self, features, labels=None):
        labels, mode = L.validate_inputs(features, labels, None, self.contrast_mode)
        if OPS.requires_cuda and not features.is_cuda:
            raise RuntimeError('DistributedSupConLoss runs on CUDA tensors only (there is no CPU fallback); '
                               'got a %s tensor' % features.device.type)
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError('DistributedSupConLoss needs an initialised torch.distributed process group; '
                               'use SupConLoss on a single GPU')
        group = self.process_group if self.process_group is not None else dist.group.WORLD
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        if len(features.shape) > 3:
            features = features.view(features.shape[0], features.shape[1], -1)
        if features.stride(2) != 1 and features.shape[2] > 1:
            features = features.contiguous()
        bsz, n_views, dim = features.shape
        dev = features.device
        labels_all = None
        mask_kind = _lib.MASK_SIMCLR
        if labels is not None:
            mask_kind = _lib.MASK_LABELS
            labels = labels.to(device=dev, dtype=torch.int64).contiguous()
            labels_all = torch.empty(world * bsz, dtype=torch.int64, device=dev)
            _all_gather_flat(labels_all, labels, group)       # sample id = rank*bsz + b
        engine = L.select_engine(self.engine, features.dtype, dim, mask_kind, world * n_views * bsz)
        plan = L.LossPlan(bsz, n_views, dim, n_seg=world, seg_index=rank, mode=mode, mask_kind=mask_kind,
                          engine=engine)
        plan.scalar_dtype = _lib.dtype_code(features.dtype)
        return _DistSupConFunction.apply(features, labels_all, plan, float(self.temperature),
                                         float(self.base_temperature), bool(self.normalize), group)
