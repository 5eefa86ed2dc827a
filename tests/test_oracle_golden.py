"""Pin the CPU oracle against vectors produced by the unmodified reference
(tests/golden/make_golden.py).  CPU only."""
import numpy as np
import pytest
import torch

from conftest import golden, golden_names
from oracle import (supcon_forward_backward, gather_glue, normalize_rows_backward,
                    syncbn_forward, syncbn_backward)
from oracle.dense_port import dense_supcon_fwd_bwd, bn_train_fwd_bwd


def _loss_args(g):
    kw = dict(temperature=float(g['temperature']), base_temperature=float(g['base_temperature']),
              contrast_mode=str(g['mode']))
    return g.get('labels'), g.get('mask'), kw


@pytest.mark.parametrize('name', golden_names('loss_'))
def test_supcon_oracle_matches_reference(name):
    g = golden(name)
    labels, mask, kw = _loss_args(g)
    loss, grad, _ = supcon_forward_backward(g['features'], labels, mask, **kw)
    f32 = g['features'].dtype == np.float32
    # a user mask is `.float()`-ed by the reference (losses.py:50) and its row sums are then
    # accumulated in fp32 (:87), so those cases carry the reference's own ~1e-7 rounding
    rtol = 2e-6 if f32 else (1e-6 if mask is not None else 1e-11)
    assert abs(loss - float(g['loss'])) <= rtol * abs(float(g['loss']))
    scale = np.abs(g['grad']).max()
    assert np.abs(grad.reshape(g['grad'].shape) - g['grad']).max() <= 2.5 * rtol * scale


@pytest.mark.parametrize('name', golden_names('loss_'))
def test_dense_port_matches_reference(name):
    g = golden(name)
    labels, mask, kw = _loss_args(g)
    f = torch.from_numpy(g['features'])
    loss, grad = dense_supcon_fwd_bwd(f, None if labels is None else torch.from_numpy(labels),
                                      None if mask is None else torch.from_numpy(mask), **kw)
    assert np.allclose(loss.numpy(), g['loss'], rtol=1e-6, atol=0)
    assert np.allclose(grad.numpy(), g['grad'], rtol=1e-5, atol=1e-7 * np.abs(g['grad']).max())


def test_oracle_error_behaviour():
    z = np.zeros((4, 8))
    with pytest.raises(ValueError, match='at least 3 dimensions'):
        supcon_forward_backward(z)
    z = np.random.randn(4, 2, 8)
    with pytest.raises(ValueError, match='Cannot define both'):
        supcon_forward_backward(z, labels=np.arange(4), mask=np.eye(4))
    with pytest.raises(ValueError, match='Num of labels'):
        supcon_forward_backward(z, labels=np.arange(3))
    with pytest.raises(ValueError, match='Unknown mode'):
        supcon_forward_backward(z, contrast_mode='two')


@pytest.mark.parametrize('name', golden_names('glue_'))
def test_gather_glue_matches_reference(name):
    g = golden(name)
    W, T = int(g['world']), float(g['temperature'])
    parts = [g[f'part{r}'] for r in range(W)]
    n_features, index = gather_glue(parts, n_views=2)
    assert np.allclose(n_features, g['n_features'], rtol=0, atol=1e-14)
    loss, dfeat, _ = supcon_forward_backward(n_features, temperature=T)
    assert abs(loss - float(g['loss'])) < 1e-11 * abs(float(g['loss']))
    D = parts[0].shape[1]
    flat = dfeat.reshape(-1, D)
    for r in range(W):
        g_hat = flat[index[r]]                        # d loss / d (normalised local rows)
        g_part = normalize_rows_backward(parts[r], g_hat)
        assert np.abs(g_part - g[f'grad{r}']).max() < 1e-11 * np.abs(g[f'grad{r}']).max()


@pytest.mark.parametrize('name', golden_names('bn_'))
def test_syncbn_oracle_matches_torch(name):
    g = golden(name)
    W = len(g['counts'])
    xs = [g[f'x{r}'] for r in range(W)]
    dys = [g[f'dy{r}'] for r in range(W)]
    f32 = xs[0].dtype == np.float32
    tol = 2e-5 if f32 else 1e-11
    ys, mean, invstd, n, rm, rv = syncbn_forward(
        xs, g['weight'], g['bias'], float(g['eps']), g['running_mean0'].astype(np.float64),
        g['running_var0'].astype(np.float64), float(g['momentum']))
    assert n == sum(int(c) * xs[0][0, 0].size for c in g['counts'])
    for r in range(W):
        assert np.abs(ys[r] - g[f'y{r}']).max() < tol * max(1.0, np.abs(g[f'y{r}']).max())
    assert np.allclose(rm, g['running_mean1'], rtol=tol, atol=tol)
    assert np.allclose(rv, g['running_var1'], rtol=tol, atol=tol)
    dxs, gws, gbs, _, _ = syncbn_backward(xs, dys, mean, invstd, g['weight'])
    for r in range(W):
        assert np.abs(dxs[r] - g[f'dx{r}']).max() < tol * max(1.0, np.abs(g[f'dx{r}']).max())
    # per-rank grad_weight/grad_bias are left un-synced by the reference; their sum is the
    # single-process gradient
    assert np.allclose(np.sum(gws, 0), g['grad_weight'], rtol=10 * tol, atol=10 * tol)
    assert np.allclose(np.sum(gbs, 0), g['grad_bias'], rtol=10 * tol, atol=10 * tol)


def test_bn_port_matches_golden():
    g = golden('bn_w2_f32')
    x = torch.from_numpy(np.concatenate([g['x0'], g['x1']]))
    dy = torch.from_numpy(np.concatenate([g['dy0'], g['dy1']]))
    rm, rv = torch.from_numpy(g['running_mean0'].copy()), torch.from_numpy(g['running_var0'].copy())
    y, dx, dw, db = bn_train_fwd_bwd(x, dy, torch.from_numpy(g['weight']), torch.from_numpy(g['bias']),
                                     rm, rv, float(g['momentum']), float(g['eps']))
    assert np.allclose(y.numpy(), np.concatenate([g['y0'], g['y1']]), atol=1e-5)
    assert np.allclose(dx.numpy(), np.concatenate([g['dx0'], g['dx1']]), atol=1e-5)
    assert np.allclose(rm.numpy(), g['running_mean1'], atol=1e-6)
