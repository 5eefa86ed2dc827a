"""numpy restatement of the reference's SupCon / NT-Xent loss.  TEST INFRASTRUCTURE ONLY
(see oracle/__init__.py): never imported by the product package.

Follows /root/reference/losses.py:17-93 (forward) and the autograd backward that PyTorch
derives from it, written here in closed form, plus the gather /
normalise glue of /root/reference/main_supcon.py:266-286 generalised from 2 to W ranks.

Everything is computed in float64 unless `dtype` says otherwise, in row blocks so that the
2N x 2N matrices are never held in full (the oracle must finish 2N=8192 in seconds).
"""
from __future__ import annotations

import numpy as np


# --------------------------------------------------------------------------- helpers
def _contrast_matrix(features: np.ndarray) -> np.ndarray:
    """losses.py:52-53 -- `cat(unbind(features, 1))`: row r = v*B + b holds features[b, v]."""
    B, V = features.shape[:2]
    f = features.reshape(B, V, -1)
    return np.ascontiguousarray(np.transpose(f, (1, 0, 2)).reshape(V * B, -1))


def _base_mask(B: int, labels, mask, dtype) -> np.ndarray:
    """losses.py:39-50 -- identity / label equality / user mask (may be asymmetric, any float)."""
    if labels is not None and mask is not None:
        raise ValueError('Cannot define both `labels` and `mask`')
    if labels is None and mask is None:
        return np.eye(B, dtype=dtype)
    if labels is not None:
        lab = np.asarray(labels).reshape(-1, 1)
        if lab.shape[0] != B:
            raise ValueError('Num of labels does not match num of features')
        return (lab == lab.T).astype(dtype)
    return np.asarray(mask).astype(np.float32).astype(dtype)     # `mask.float()` :50 rounds to fp32


def supcon_row_stats(features, labels=None, mask=None, temperature=0.07,
                     contrast_mode='all', block=512, dtype=np.float64):
    """Per-anchor quantities of the forward pass (what the CUDA forward kernel emits).

    Returns (lse[A], possum[A], npos[A]) with
      lse_i    = log sum_{j != i} exp(s_ij)                 losses.py:83-84 (shift-free form)
      possum_i = sum_j M_ij s_ij,  npos_i = sum_j M_ij      losses.py:87
    where s = anchor . contrast^T / T (losses.py:64-66) and M is the tiled base mask with the
    self-contrast column removed (losses.py:72-80).
    """
    features = np.asarray(features)
    if features.ndim < 3:
        raise ValueError('`features` needs to be [bsz, n_views, ...],'
                         'at least 3 dimensions are required')
    B, V = features.shape[:2]
    K = _contrast_matrix(features).astype(dtype)
    m = _base_mask(B, labels, mask, dtype)
    if contrast_mode == 'one':
        A = B
    elif contrast_mode == 'all':
        A = V * B
    else:
        raise ValueError('Unknown mode: {}'.format(contrast_mode))
    C = V * B
    lse = np.empty(A, dtype)
    possum = np.empty(A, dtype)
    npos = np.empty(A, dtype)
    cols = np.arange(C)
    for r0 in range(0, A, block):
        r1 = min(A, r0 + block)
        s = (K[r0:r1] @ K.T) / dtype(temperature)                 # losses.py:64-66
        rows = np.arange(r0, r1)
        notself = (cols[None, :] != rows[:, None])                 # losses.py:74-79
        M = m[rows % B][:, cols % B] * notself                     # losses.py:72,80
        smax = s.max(axis=1, keepdims=True)                        # losses.py:68 (incl. self)
        e = np.exp(s - smax) * notself                             # losses.py:83
        lse[r0:r1] = np.log(e.sum(axis=1)) + smax[:, 0]            # losses.py:84
        possum[r0:r1] = (M * s).sum(axis=1)
        npos[r0:r1] = M.sum(axis=1)
    return lse, possum, npos


def supcon_forward_backward(features, labels=None, mask=None, temperature=0.07,
                            contrast_mode='all', base_temperature=0.07,
                            block=512, dtype=np.float64, grad_out=1.0):
    """Loss scalar and d loss / d features, same argument meaning as the reference's
    `SupConLoss(temperature, contrast_mode, base_temperature).forward(features, labels, mask)`
    (losses.py:10-17) followed by `loss.backward()`.

    Closed form:
      loss_i = -(T/Tb) (possum_i/npos_i - lse_i),  L = mean_i loss_i       losses.py:87-91
      G_ij   = (T/Tb)/A (P_ij - M_ij/npos_i),  P_ij = exp(s_ij - lse_i)[i != j]
      dAnchor = G K / T,  dContrast = G^T Anchor / T,  d features[b, v] = dK[v*B + b]
    """
    features = np.asarray(features)
    lse, possum, npos = supcon_row_stats(features, labels, mask, temperature,
                                         contrast_mode, block, dtype)
    B, V = features.shape[:2]
    K = _contrast_matrix(features).astype(dtype)
    m = _base_mask(B, labels, mask, dtype)
    A = lse.shape[0]
    C = V * B
    T = dtype(temperature)
    scale = dtype(temperature) / dtype(base_temperature)
    with np.errstate(divide='ignore', invalid='ignore'):
        loss_rows = -scale * (possum / npos - lse)                 # losses.py:87,90 (0/0 -> nan)
    loss = loss_rows.mean()                                        # losses.py:91
    dK = np.zeros_like(K)
    cols = np.arange(C)
    coef = dtype(grad_out) * scale / dtype(A)
    for r0 in range(0, A, block):
        r1 = min(A, r0 + block)
        rows = np.arange(r0, r1)
        s = (K[r0:r1] @ K.T) / T
        notself = (cols[None, :] != rows[:, None])
        M = m[rows % B][:, cols % B] * notself
        P = np.exp(s - lse[r0:r1, None]) * notself
        with np.errstate(divide='ignore', invalid='ignore'):
            G = coef * (P - M / npos[r0:r1, None])
        dK[r0:r1] += (G @ K) / T                                   # d anchor rows
        dK += (G.T @ K[r0:r1]) / T                                 # d contrast rows
    dfeat = np.transpose(dK.reshape(V, B, -1), (1, 0, 2)).reshape(features.shape)
    return loss, dfeat, loss_rows


# ----------------------------------------------------------------- normalise + gather glue
def normalize_rows(z, eps=1e-12):
    """`F.normalize(features, dim=1)` of main_supcon.py:283: z / max(||z||, eps)."""
    z = np.asarray(z, dtype=np.float64)
    n = np.maximum(np.sqrt((z * z).sum(axis=-1, keepdims=True)), eps)
    return z / n


def normalize_rows_backward(z, g, eps=1e-12):
    """Backward of `F.normalize` (main_supcon.py:283): dz = (g - zh (zh . g)) / max(||z||, eps)."""
    z = np.asarray(z, dtype=np.float64)
    g = np.asarray(g, dtype=np.float64)
    n = np.maximum(np.sqrt((z * z).sum(axis=-1, keepdims=True)), eps)
    zh = z / n
    return (g - zh * (zh * g).sum(axis=-1, keepdims=True)) / n


def gather_glue(part_features_per_rank, n_views=2):
    """W-rank restatement of main_supcon.py:266-286.

    Each rank r holds `part_features[r]` of shape [n_views*bsz, D] (view-major: all view-0
    embeddings of its bsz samples, then view 1, ...; main_supcon.py:256,266).  The reference
    all-gathers them (:268-269), re-orders to [v0 of every rank ; v1 of every rank] (:276-279,
    hard-wired to W=2 there), L2-normalises (:283) and stacks the views (:285-286) into
    n_features [W*bsz, n_views, D], which every rank feeds to the criterion.

    Returns (n_features, index) where index[r] is an int array [n_views*bsz] with the flat
    position (b_global*n_views + v) of each of rank r's local rows inside n_features, so a
    test can cut rank r's gradient slice out of d loss / d n_features.
    """
    W = len(part_features_per_rank)
    parts = [np.asarray(p, dtype=np.float64) for p in part_features_per_rank]
    D = parts[0].shape[1]
    bsz = parts[0].shape[0] // n_views
    views = []
    for v in range(n_views):
        views.append(np.concatenate([p[v * bsz:(v + 1) * bsz] for p in parts], axis=0))
    feats = np.concatenate(views, axis=0)                           # [V*W*bsz, D]  (:276-279)
    n_fea = normalize_rows(feats)                                   # (:283)
    n_features = np.stack(np.split(n_fea, n_views, axis=0), axis=1)  # [W*bsz, V, D] (:285-286)
    index = []
    for r in range(W):
        idx = np.empty(n_views * bsz, dtype=np.int64)
        for v in range(n_views):
            b_global = r * bsz + np.arange(bsz)
            idx[v * bsz:(v + 1) * bsz] = b_global * n_views + v
        index.append(idx)
    assert n_features.shape == (W * bsz, n_views, D)
    return n_features, index
