"""Parity of the CUDA loss (through the Python boundary and the C ABI) with the oracle and the
golden vectors of the unmodified reference.  Tolerances are the north_star's: <= 1e-4 relative
in the fp32 / fp32-accumulate modes ('fp32', 'bf16x3'), <= 2e-3 in 'bf16' (bf16 operands; the
oracle is fed the same bf16-rounded embeddings)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from conftest import golden, golden_names
from oracle import supcon_forward_backward, normalize_rows, normalize_rows_backward
from simclr_pytorch_distributed_b200 import SupConLoss
from simclr_pytorch_distributed_b200 import losses as L, _lib

pytestmark = pytest.mark.gpu

TOL = {'fp32': 1e-4, 'bf16x3': 1e-4, 'bf16': 2e-3}


def _available(engine):
    lib = _lib.load()
    if engine == 'fp32':
        return True
    try:
        f = torch.zeros(4, 2, 64, device='cuda')
        f[:, :, 0] = 1
        SupConLoss(engine=engine)(f)
        return True
    except RuntimeError as e:
        if 'not built yet' in str(e):
            return False
        raise


def engines():
    return ['fp32', 'bf16x3', 'bf16']


def run_ours(feat_np, labels, mask, engine, dtype=torch.float32, **kw):
    f = torch.from_numpy(np.asarray(feat_np)).to('cuda', dtype).requires_grad_(True)
    lab = None if labels is None else torch.from_numpy(np.asarray(labels)).cuda()
    m = None if mask is None else torch.from_numpy(np.asarray(mask)).cuda()
    loss = SupConLoss(engine=engine, **kw)(f, labels=lab, mask=m)
    loss.backward()
    torch.cuda.synchronize()
    return float(loss), f.grad.float().cpu().numpy(), f.detach().float().cpu().numpy()


def assert_close(loss, grad, ref_loss, ref_grad, tol, what=''):
    assert np.isfinite(loss), what
    assert abs(loss - ref_loss) <= tol * abs(ref_loss) + 1e-6, '%s loss %r vs %r' % (what, loss, ref_loss)
    num = np.linalg.norm((grad - ref_grad).ravel())
    den = np.linalg.norm(ref_grad.ravel())
    assert num <= tol * den + 1e-9, '%s grad rel-F %.3e' % (what, num / den)
    assert np.abs(grad - ref_grad).max() <= 5 * tol * np.abs(ref_grad).max() + 1e-9, what


@pytest.mark.parametrize('engine', engines())
@pytest.mark.parametrize('name', golden_names('loss_'))
def test_golden_vectors(name, engine):
    if not _available(engine):
        pytest.skip('engine not built')
    g = golden(name)
    labels, mask = g.get('labels'), g.get('mask')
    kw = dict(temperature=float(g['temperature']), base_temperature=float(g['base_temperature']),
              contrast_mode=str(g['mode']))
    if engine != 'fp32' and mask is not None:
        with pytest.raises(ValueError, match='does not take this problem'):
            run_ours(g['features'], labels, mask, engine, **kw)
        return
    if engine == 'bf16':
        # operands are bf16: feed the oracle the same rounded embeddings
        loss, grad, f_used = run_ours(g['features'], labels, mask, engine, dtype=torch.bfloat16, **kw)
        ref_loss, ref_grad, _ = supcon_forward_backward(f_used.astype(np.float64), labels, mask, **kw)
        ref_grad = ref_grad.reshape(grad.shape)
        # the returned tensors are bf16 themselves: allow their rounding on top
        assert abs(loss - ref_loss) <= (TOL[engine] + 2 ** -8) * abs(ref_loss)
        num = np.linalg.norm((grad - ref_grad).ravel()) / np.linalg.norm(ref_grad.ravel())
        assert num <= TOL[engine] + 2 ** -8, num
    else:
        loss, grad, _ = run_ours(g['features'].astype(np.float32), labels, mask, engine, **kw)
        assert_close(loss, grad, float(g['loss']), g['grad'].reshape(grad.shape).astype(np.float64),
                     TOL[engine], name)


SHAPES = [(2, 2, 8), (37, 2, 50), (130, 3, 128), (64, 1, 64), (300, 2, 20), (129, 2, 256), (96, 4, 72)]


@pytest.mark.parametrize('engine', engines())
@pytest.mark.parametrize('mode', ['all', 'one'])
@pytest.mark.parametrize('shape', SHAPES)
def test_ragged_shapes_simclr_and_labels(shape, mode, engine):
    if not _available(engine):
        pytest.skip('engine not built')
    B, V, D = shape
    if engine != 'fp32' and not L.tc_engine_supported(D, _lib.MASK_SIMCLR, B * V):
        # shapes outside the tensor-core engine are refused loudly, never re-routed silently
        with pytest.raises(ValueError, match='does not take this problem'):
            run_ours(np.zeros(shape, np.float32), None, None, engine)
        return
    rng = np.random.default_rng(B * 1000 + V * 10 + D)
    z = rng.standard_normal((B, V, D)) + 0.5 * rng.standard_normal((1, 1, D))
    z = (z / np.linalg.norm(z, axis=-1, keepdims=True)).astype(np.float32)
    labels = rng.integers(0, max(2, B // 4), size=B)
    for lab in ([None, labels] if V > 1 else [labels]):
        if V == 1 and len(np.unique(labels)) == B:
            continue
        T = 0.1 if lab is not None else 0.5
        # every anchor needs a positive, else the reference itself returns nan (losses.py:87)
        if lab is not None and V == 1:
            lab = np.arange(B) % max(1, B // 2)
        dtype = torch.bfloat16 if engine == 'bf16' else torch.float32
        loss, grad, used = run_ours(z, lab, None, engine, dtype=dtype, temperature=T, contrast_mode=mode)
        ref_loss, ref_grad, _ = supcon_forward_backward(used.astype(np.float64), lab, None,
                                                        temperature=T, contrast_mode=mode)
        tol = TOL[engine] + (2 ** -8 if engine == 'bf16' else 0)
        assert abs(loss - ref_loss) <= tol * abs(ref_loss) + 1e-6, (loss, ref_loss)
        rel = np.linalg.norm((grad - ref_grad).ravel()) / np.linalg.norm(ref_grad.ravel())
        assert rel <= tol, rel


def test_explicit_masks_fp32():
    rng = np.random.default_rng(5)
    B, V, D = 45, 2, 33
    z = rng.standard_normal((B, V, D)).astype(np.float32) * 0.4
    m = (rng.random((B, B)) < 0.3) * rng.random((B, B)) + np.eye(B)
    for mode in ('all', 'one'):
        loss, grad, _ = run_ours(z, None, m.astype(np.float32), 'fp32', temperature=0.3, contrast_mode=mode)
        ref_loss, ref_grad, _ = supcon_forward_backward(z.astype(np.float64), None, m, temperature=0.3,
                                                        contrast_mode=mode)
        assert_close(loss, grad, ref_loss, ref_grad, 1e-4, mode)


def test_auto_engine_takes_explicit_mask_and_large_dim():
    rng = np.random.default_rng(6)
    z = rng.standard_normal((16, 2, 300)).astype(np.float32)
    z /= np.linalg.norm(z, axis=-1, keepdims=True)
    loss, grad, _ = run_ours(z, None, None, 'auto', temperature=0.5)
    ref_loss, ref_grad, _ = supcon_forward_backward(z.astype(np.float64), temperature=0.5)
    assert_close(loss, grad, ref_loss, ref_grad, 1e-4)


@pytest.mark.parametrize('engine', engines())
def test_fused_normalize_matches_glue(engine):
    """normalize=True fuses main_supcon.py:283 (F.normalize) and its backward."""
    if not _available(engine):
        pytest.skip('engine not built')
    rng = np.random.default_rng(7)
    raw = (rng.standard_normal((64, 2, 128)) * 3.0).astype(np.float32)
    f = torch.from_numpy(raw).cuda().requires_grad_(True)
    loss = SupConLoss(temperature=0.5, engine=engine, normalize=True)(f)
    loss.backward()
    zh = normalize_rows(raw)
    ref_loss, g_hat, _ = supcon_forward_backward(zh, temperature=0.5)
    ref_grad = normalize_rows_backward(raw, g_hat)
    tol = 5e-3 if engine == 'bf16' else 1e-4      # bf16: the normalised rows are rounded to bf16
    assert abs(float(loss) - ref_loss) <= tol * abs(ref_loss)
    rel = np.linalg.norm(f.grad.cpu().numpy().ravel() - ref_grad.ravel()) / np.linalg.norm(ref_grad.ravel())
    assert rel <= tol, rel


def test_value_errors_on_gpu_tensors():
    f = torch.zeros(4, 2, 8, device='cuda')
    with pytest.raises(ValueError, match='at least 3 dimensions'):
        SupConLoss()(torch.zeros(4, 8, device='cuda'))
    with pytest.raises(ValueError, match='Cannot define both'):
        SupConLoss()(f, labels=torch.arange(4), mask=torch.eye(4))
    with pytest.raises(ValueError, match='Num of labels'):
        SupConLoss()(f, labels=torch.arange(3))
    with pytest.raises(ValueError, match='Unknown mode'):
        SupConLoss(contrast_mode='x')(f)


def test_no_positive_gives_nan_like_reference():
    z = F.normalize(torch.randn(8, 1, 16, device='cuda'), dim=-1)
    assert torch.isnan(SupConLoss(engine='fp32')(z))            # losses.py:87: 0/0


def test_more_than_three_dims_and_grad_scaling():
    rng = np.random.default_rng(9)
    z = rng.standard_normal((12, 2, 2, 3, 4)).astype(np.float32)
    f = torch.from_numpy(z).cuda().requires_grad_(True)
    loss = SupConLoss(temperature=0.5, engine='fp32')(f)
    (3.0 * loss).backward()                                     # upstream gradient != 1
    ref_loss, ref_grad, _ = supcon_forward_backward(z.astype(np.float64), temperature=0.5)
    assert f.grad.shape == f.shape
    assert_close(float(loss), f.grad.cpu().numpy() / 3.0, ref_loss, ref_grad.reshape(z.shape), 1e-4)


@pytest.mark.parametrize('engine', engines())
def test_full_size_cfg4_against_oracle(engine):
    """BASELINE config 4 on one GPU: 2N = 8192, D = 128, T = 0.5 (oracle in row blocks, fp64)."""
    if not _available(engine):
        pytest.skip('engine not built')
    g = torch.Generator().manual_seed(1234)
    z = F.normalize(torch.randn(4096, 2, 128, generator=g) + 0.5 * torch.randn(1, 1, 128, generator=g), dim=-1)
    dtype = torch.bfloat16 if engine == 'bf16' else torch.float32
    loss, grad, used = run_ours(z.numpy(), None, None, engine, dtype=dtype, temperature=0.5)
    ref_loss, ref_grad, _ = supcon_forward_backward(used.astype(np.float64), temperature=0.5, block=1024)
    tol = TOL[engine] + (2 ** -8 if engine == 'bf16' else 0)
    assert abs(loss - ref_loss) <= tol * abs(ref_loss)
    rel = np.linalg.norm((grad - ref_grad).ravel()) / np.linalg.norm(ref_grad.ravel())
    assert rel <= tol, rel


@pytest.mark.parametrize('engine', engines())
def test_size_independent_properties_full_size(engine):
    """Properties that need no oracle at 2N = 8192: permutation invariance of the loss under a
    consistent re-ordering of the samples, and agreement of the analytic gradient with a central
    difference of our own (fp32-engine) forward along the gradient direction."""
    if not _available(engine):
        pytest.skip('engine not built')
    torch.manual_seed(3)
    z = F.normalize(torch.randn(4096, 2, 128, device='cuda'), dim=-1)
    crit = SupConLoss(temperature=0.5, engine=engine)
    perm = torch.randperm(4096, device='cuda')
    l0, l1 = float(crit(z)), float(crit(z[perm]))
    assert abs(l0 - l1) <= 2e-6 * abs(l0)
    f = z.clone().requires_grad_(True)
    crit(f).backward()
    d = f.grad.float() / f.grad.float().norm()       # steepest direction: the signal is ||grad||
    eps = 1e-2
    exact = SupConLoss(temperature=0.5, engine='fp32')
    num = (float(exact(z + eps * d)) - float(exact(z - eps * d))) / (2 * eps)
    ana = float((f.grad * d).sum())
    assert abs(num - ana) <= 2e-2 * abs(ana) + 1e-5, (num, ana)
