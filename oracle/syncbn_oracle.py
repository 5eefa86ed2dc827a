"""numpy restatement of the SyncBatchNorm arithmetic the reference runs once `--syncBN`
swaps `nn.SyncBatchNorm` into the network (/root/reference/main_supcon.py:222-224).
TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

The arithmetic itself is third-party: module `torch` (reference README pins PyTorch 1.2.0;
this image has 2.11.0+cu128).  Restated from
  torch/nn/modules/_functions.py:10-124   (forward:  stats -> all_gather -> merge -> elemt)
  torch/nn/modules/_functions.py:127-209  (backward: reduce -> all_reduce -> elemt)
  torch/include/ATen/native/cuda/Normalization.cuh:458-502 (merge with counts + running stats)
Those ATen ops are CUDA-only, so parity is pinned on `torch.nn.BatchNorm2d` over the
rank-concatenated batch (mathematically the same thing), see tests/golden/make_golden.py.

Arrays are logical [N, C, *spatial]; the reduction runs over every axis except 1.
"""
from __future__ import annotations

import numpy as np


def _red_axes(x):
    return tuple(a for a in range(x.ndim) if a != 1)


def _bc(v, x):
    shape = [1] * x.ndim
    shape[1] = -1
    return v.reshape(shape)


def bn_local_stats(x, eps=1e-5, dtype=np.float64):
    """`torch.batch_norm_stats` (_functions.py:39): per-channel mean and 1/sqrt(biased var + eps)
    of one rank's slice, plus the per-channel element count (_functions.py:41-46)."""
    x = np.asarray(x).astype(dtype)
    ax = _red_axes(x)
    mean = x.mean(axis=ax)
    var = ((x - _bc(mean, x)) ** 2).mean(axis=ax)
    invstd = 1.0 / np.sqrt(var + eps)
    count = x.size // x.shape[1]
    return mean, invstd, count


def bn_merge_stats(means, invstds, counts, eps=1e-5, running_mean=None, running_var=None,
                   momentum=0.1):
    """`torch.batch_norm_gather_stats_with_counts` (_functions.py:106-115;
    Normalization.cuh:458-502): pairwise (Chan) merge of the ranks' (mean, var, n) in rank order,
    var_r recovered as 1/invstd_r^2 - eps; running_var uses the unbiased global variance."""
    mean = np.zeros_like(means[0])
    m2 = np.zeros_like(means[0])
    n = 0.0
    for mu_r, is_r, n_r in zip(means, invstds, counts):
        if n_r < 1:          # ranks with empty input are filtered out (_functions.py:97-100)
            continue
        var_r = 1.0 / (is_r * is_r) - eps
        delta = mu_r - mean
        tot = n + n_r
        mean = mean + delta * (n_r / tot)
        m2 = m2 + var_r * n_r + delta * delta * (n * n_r / tot)
        n = tot
    invstd = 1.0 / np.sqrt(m2 / n + eps)
    new_rm, new_rv = running_mean, running_var
    if running_mean is not None:
        new_rm = (1.0 - momentum) * running_mean + momentum * mean
    if running_var is not None:
        new_rv = (1.0 - momentum) * running_var + momentum * (m2 / (n - 1.0))
    return mean, invstd, n, new_rm, new_rv


def syncbn_forward(xs, weight=None, bias=None, eps=1e-5, running_mean=None, running_var=None,
                   momentum=0.1, dtype=np.float64):
    """Training-mode forward over W ranks (xs = list of per-rank inputs).

    Returns (ys, mean, invstd, n_total, new_running_mean, new_running_var);
    y = (x - mean) * invstd * weight + bias  (`torch.batch_norm_elemt`, _functions.py:122)."""
    stats = [bn_local_stats(x, eps, dtype) for x in xs]
    mean, invstd, n, rm, rv = bn_merge_stats([s[0] for s in stats], [s[1] for s in stats],
                                             [s[2] for s in stats], eps,
                                             running_mean, running_var, momentum)
    ys = []
    for x in xs:
        x = np.asarray(x).astype(dtype)
        y = (x - _bc(mean, x)) * _bc(invstd, x)
        if weight is not None:
            y = y * _bc(np.asarray(weight, dtype), x)
        if bias is not None:
            y = y + _bc(np.asarray(bias, dtype), x)
        ys.append(y)
    return ys, mean, invstd, n, rm, rv


def syncbn_backward(xs, dys, mean, invstd, weight=None, dtype=np.float64):
    """Training-mode backward over W ranks.

    `torch.batch_norm_backward_reduce` (_functions.py:139-153): per rank sum_dy, sum_dy_xmu,
    grad_weight = sum_dy_xmu * invstd, grad_bias = sum_dy (left un-synced, :180-181);
    all_reduce(SUM) of [sum_dy, sum_dy_xmu] (:155-165);
    `torch.batch_norm_backward_elemt` (:170-179):
       dx = (dy - sum_dy/n - (x - mean) * invstd^2 * sum_dy_xmu/n) * weight * invstd.
    Returns (dxs, grad_weights_per_rank, grad_biases_per_rank, sum_dy_global, sum_dy_xmu_global)."""
    sum_dy_r, sum_dy_xmu_r, n = [], [], 0
    for x, dy in zip(xs, dys):
        x = np.asarray(x).astype(dtype)
        dy = np.asarray(dy).astype(dtype)
        ax = _red_axes(x)
        sum_dy_r.append(dy.sum(axis=ax))
        sum_dy_xmu_r.append((dy * (x - _bc(mean, x))).sum(axis=ax))
        n += x.size // x.shape[1]
    gws = [s * invstd for s in sum_dy_xmu_r]
    gbs = [s.copy() for s in sum_dy_r]
    sum_dy = np.sum(sum_dy_r, axis=0)
    sum_dy_xmu = np.sum(sum_dy_xmu_r, axis=0)
    w = np.ones_like(mean) if weight is None else np.asarray(weight, dtype)
    dxs = []
    for x, dy in zip(xs, dys):
        x = np.asarray(x).astype(dtype)
        dy = np.asarray(dy).astype(dtype)
        k = _bc(invstd * invstd * sum_dy_xmu / n, x)
        dx = (dy - _bc(sum_dy / n, x) - (x - _bc(mean, x)) * k) * _bc(w * invstd, x)
        dxs.append(dx)
    return dxs, gws, gbs, sum_dy, sum_dy_xmu
