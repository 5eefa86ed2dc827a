"""Parity cases the north_star names explicitly: BASELINE configs[4] at full size (SupCon with labels, batch 1024,
100 classes, T = 0.07, contrast_mode all vs one, /root/reference/losses.py:44-48,54-59,87) and the bf16 engine held
to a FLAT 2e-3 with fp32 features in and out (no allowance for a bf16 output).  The oracle is fed the embeddings the
engine actually multiplies (fp32 features rounded to bf16 once)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import supcon_forward_backward, normalize_rows, normalize_rows_backward
from simclr_pytorch_distributed_b200 import SupConLoss
from simclr_pytorch_distributed_b200.losses import GraphCache

pytestmark = pytest.mark.gpu


def rel_f(a, b):
    return np.linalg.norm((a - b).ravel()) / np.linalg.norm(b.ravel())


def run(z, labels, engine, **kw):
    f = torch.from_numpy(z).cuda().requires_grad_(True)
    lab = None if labels is None else torch.from_numpy(labels).cuda()
    loss = SupConLoss(engine=engine, **kw)(f, labels=lab)
    loss.backward()
    torch.cuda.synchronize()
    return float(loss), f.grad.cpu().numpy().astype(np.float64)


def unit_rows(rng, B, V, D, common=0.5):
    z = rng.standard_normal((B, V, D)) + common * rng.standard_normal((1, 1, D))
    return (z / np.linalg.norm(z, axis=-1, keepdims=True)).astype(np.float32)


@pytest.mark.parametrize('engine', ['fp32', 'bf16x3', 'bf16'])
@pytest.mark.parametrize('mode', ['all', 'one'])
def test_cfg5_full_size_labels(mode, engine):
    """B = 1024, V = 2, D = 128, 100 classes, T = base_T = 0.07."""
    rng = np.random.default_rng(2024)
    z = unit_rows(rng, 1024, 2, 128)
    labels = rng.integers(0, 100, size=1024)
    kw = dict(temperature=0.07, base_temperature=0.07, contrast_mode=mode)
    loss, grad = run(z, labels, engine, **kw)
    used = z if engine != 'bf16' else torch.from_numpy(z).to(torch.bfloat16).float().numpy()
    ref_loss, ref_grad, _ = supcon_forward_backward(used.astype(np.float64), labels, None, **kw)
    tol = 1e-4 if engine != 'bf16' else 2e-3
    assert abs(loss - ref_loss) <= tol * abs(ref_loss), (loss, ref_loss)
    assert rel_f(grad, ref_grad.reshape(grad.shape)) <= tol, rel_f(grad, ref_grad.reshape(grad.shape))


CASES = [('simclr', 'all', 0.5), ('simclr', 'one', 0.5), ('labels', 'all', 0.1), ('labels', 'one', 0.1)]


@pytest.mark.parametrize('kind,mode,T', CASES)
@pytest.mark.parametrize('shape', [(512, 2, 128), (300, 2, 64), (130, 3, 128)])
def test_bf16_engine_flat_2e3_with_fp32_io(shape, kind, mode, T):
    B, V, D = shape
    rng = np.random.default_rng(B + V + D)
    z = unit_rows(rng, B, V, D)
    labels = rng.integers(0, max(2, B // 8), size=B) if kind == 'labels' else None
    kw = dict(temperature=T, contrast_mode=mode)
    loss, grad = run(z, labels, 'bf16', **kw)
    used = torch.from_numpy(z).to(torch.bfloat16).double().numpy()
    ref_loss, ref_grad, _ = supcon_forward_backward(used, labels, None, **kw)
    assert abs(loss - ref_loss) <= 2e-3 * abs(ref_loss), (loss, ref_loss)
    r = rel_f(grad, ref_grad.reshape(grad.shape))
    assert r <= 2e-3, r


def test_bf16_engine_flat_2e3_fused_normalize():
    rng = np.random.default_rng(77)
    raw = (rng.standard_normal((512, 2, 128)) * 3.0).astype(np.float32)
    f = torch.from_numpy(raw).cuda().requires_grad_(True)
    loss = SupConLoss(temperature=0.5, engine='bf16', normalize=True)(f)
    loss.backward()
    # the engine multiplies the normalised rows rounded to bf16: the oracle gets exactly those, and the
    # normalisation backward is taken around the un-rounded rows (what the kernel does)
    zh = normalize_rows(raw)
    used = torch.from_numpy(zh).to(torch.bfloat16).double().numpy()
    ref_loss, g_hat, _ = supcon_forward_backward(used, temperature=0.5)
    ref_grad = normalize_rows_backward(raw, g_hat)
    assert abs(float(loss) - ref_loss) <= 2e-3 * abs(ref_loss)
    r = rel_f(f.grad.cpu().numpy().astype(np.float64), ref_grad.reshape(raw.shape))
    assert r <= 2e-3, r


def test_float_labels_are_not_truncated():
    """losses.py:48 compares the raw labels: 1.2 and 1.7 are different classes."""
    rng = np.random.default_rng(5)
    z = unit_rows(rng, 64, 2, 32)
    lab_f = np.tile(np.array([1.2, 1.7, 2.0, -0.0]), 16)
    lab_i = np.tile(np.array([0, 1, 2, 3]), 16)
    f = torch.from_numpy(z).cuda()
    a = SupConLoss(engine='fp32')(f, labels=torch.from_numpy(lab_f).cuda())
    b = SupConLoss(engine='fp32')(f, labels=torch.from_numpy(lab_i).cuda())
    assert float(a) == float(b)


def test_backward_general_geometry_matches_direct(monkeypatch):
    """The two backward paths (rows finished inside the kernel by a CTA pair / per-slot partials + reduce kernel)
    are the same arithmetic up to the order of two additions."""
    import os
    import subprocess
    import sys
    code = ("import torch, numpy as np, torch.nn.functional as F\n"
            "from simclr_pytorch_distributed_b200 import SupConLoss\n"
            "g = torch.Generator().manual_seed(3)\n"
            "z = F.normalize(torch.randn(4096, 2, 128, generator=g), dim=-1).to(torch.bfloat16).cuda().requires_grad_(True)\n"
            "l = SupConLoss(temperature=0.5, engine='bf16')(z); l.backward(); torch.cuda.synchronize()\n"
            "np.save(__import__('sys').argv[1], z.grad.float().cpu().numpy()); print(float(l))\n")
    outs = []
    for direct in ('1', '0'):
        path = '/tmp/scb_bwd_%s.npy' % direct
        env = dict(os.environ, SCB_BWD_DIRECT=direct)
        r = subprocess.run([sys.executable, '-c', code, path], env=env, capture_output=True, text=True, timeout=300,
                           cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append((float(r.stdout.strip().splitlines()[-1]), np.load(path)))
    assert outs[0][0] == outs[1][0]
    assert rel_f(outs[0][1].astype(np.float64), outs[1][1].astype(np.float64)) <= 2.0 ** -8


def test_auto_engine_is_the_split_bf16_engine_for_fp32_features():
    """The literal two-line change of INTEGRATION.md (fp32 features, engine 'auto', as main_supcon.py:283-290 calls
    it) lands on the tensor cores at the reference's precision: raw fp32 rows, fused normalisation, 1e-4."""
    from simclr_pytorch_distributed_b200 import losses as L, _lib
    assert L.select_engine('auto', torch.float32, 128, _lib.MASK_SIMCLR, 8192) == _lib.ENGINE_BF16X3
    assert L.select_engine('auto', torch.bfloat16, 128, _lib.MASK_SIMCLR, 8192) == _lib.ENGINE_BF16
    assert L.select_engine('auto', torch.float32, 300, _lib.MASK_SIMCLR, 8192) == _lib.ENGINE_FP32
    rng = np.random.default_rng(31)
    raw = (rng.standard_normal((1024, 2, 128)) * 2.0 + 0.3).astype(np.float32)
    f = torch.from_numpy(raw).cuda().requires_grad_(True)
    loss = SupConLoss(temperature=0.5, normalize=True)(f)
    loss.backward()
    ref_loss, g_hat, _ = supcon_forward_backward(normalize_rows(raw), temperature=0.5)
    ref_grad = normalize_rows_backward(raw, g_hat)
    assert abs(float(loss) - ref_loss) <= 1e-4 * abs(ref_loss)
    assert rel_f(f.grad.cpu().numpy().astype(np.float64), ref_grad.reshape(raw.shape)) <= 1e-4


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_feature_norm_side_outputs_and_regularisers(dtype):
    """main_supcon.py:298-317 restated with torch autograd: norm mean / variance, the moving average, loss_sec and
    loss_l2reg, and the gradient of `loss + w_sec * loss_sec + w_l2 * loss_l2reg` with respect to the raw features."""
    torch.manual_seed(21)
    B, V, D = 256, 2, 128
    momentum, w_sec, w_l2 = 0.1, 0.3, 0.05
    crit = SupConLoss(temperature=0.5, normalize=True, track_norms=True, norm_momentum=momentum)
    record = None
    for step in range(3):
        raw = (torch.randn(B, V, D, device='cuda') * (2.0 + step) + 0.1).to(dtype)
        f = raw.clone().requires_grad_(True)
        loss = crit(f, sec_weight=w_sec, l2reg_weight=w_l2)
        loss.backward()
        r = raw.double().requires_grad_(True)
        norms = torch.norm(r.view(B * V, D), p=2, dim=1)                       # :298
        norm_mean = norms.mean()
        norm_var = ((norms - norm_mean) ** 2).mean()
        record = norm_mean.detach() if record is None else (1 - momentum) * record + momentum * norm_mean.detach()
        loss_sec = ((norms - record) ** 2).mean()
        loss_l2 = (norms ** 2).mean()
        from oracle.dense_port import dense_supcon_loss
        base = dense_supcon_loss(F.normalize(r, dim=-1), temperature=0.5)
        ref = base + w_sec * loss_sec + w_l2 * loss_l2
        ref.backward()
        t = crit.norm_terms.double()
        tol = 1e-5 if dtype == torch.float32 else 1e-5        # the terms come from fp32 norms of the given values
        for got, want in zip(t[:4], (norm_mean, norm_var, loss_sec, loss_l2)):
            assert abs(float(got) - float(want)) <= tol * max(1.0, abs(float(want))) + 1e-4 * abs(float(want))
        assert abs(float(crit.record_norm_mean[0]) - float(record)) <= 1e-5 * float(record)
        ltol = 1e-4 if dtype == torch.float32 else 2e-3 + 2 ** -8
        assert abs(float(loss) - float(ref)) <= ltol * abs(float(ref))
        gtol = 1e-4 if dtype == torch.float32 else 2e-3 + 2 ** -7
        assert rel_f(f.grad.double().cpu().numpy(), r.grad.cpu().numpy()) <= gtol


def test_graph_mode_matches_eager():
    """SupConLoss(graph=True): forward+backward replayed from a CUDA graph -- same loss and gradient as the eager
    module, for recurring input buffers (no copy) and for fresh tensors (copied into the module's buffer), with and
    without labels, and an upstream gradient different from 1."""
    torch.manual_seed(8)
    eager = SupConLoss(temperature=0.5, engine='bf16')
    graphed = SupConLoss(temperature=0.5, engine='bf16', graph=True)
    bufs = [torch.empty(512, 2, 128, dtype=torch.bfloat16, device='cuda') for _ in range(2)]
    for step in range(7):
        z = F.normalize(torch.randn(512, 2, 128, device='cuda'), dim=-1).to(torch.bfloat16)
        if step < 4:
            bufs[step % 2].copy_(z)
            src = bufs[step % 2]                 # recurring addresses
        else:
            src = z.clone()                      # fresh tensors: more addresses than the module binds graphs to
        labels = torch.randint(0, 20, (512,), device='cuda') if step % 3 == 2 else None
        a = src.detach().clone().requires_grad_(True)
        b = src.detach().requires_grad_(True)
        la = eager(a, labels=labels)
        (2.0 * la).backward()
        lb = graphed(b, labels=labels)
        (2.0 * lb).backward()
        assert float(la) == float(lb)
        assert torch.equal(a.grad, b.grad)
    with pytest.raises(RuntimeError, match='followed by another forward'):
        b1 = bufs[0].detach().requires_grad_(True)
        l1 = graphed(b1)
        graphed(bufs[0].detach().requires_grad_(True))
        l1.backward()
    # a training loop's features are a fresh allocation every step; the allocator tends to hand out the same address
    # again, and the graph bound to it must then see the NEW contents (the entry does not keep the old tensor alive)
    loop = SupConLoss(temperature=0.5, engine='bf16', graph=True)
    addresses = set()
    for step in range(6):
        z = F.normalize(torch.randn(512, 2, 128, device='cuda'), dim=-1).to(torch.bfloat16).requires_grad_(True)
        addresses.add(z.data_ptr())
        a = z.detach().clone().requires_grad_(True)
        la, lb = eager(a), loop(z)
        la.backward()
        lb.backward()
        assert float(la) == float(lb)
        assert torch.equal(a.grad, z.grad)
        del z, a, la, lb
    bound = [k for per_shape in loop._graph_cache.graphs.values() for k in per_shape if k != 'own']
    assert set(bound) <= addresses and len(bound) <= GraphCache.MAX_ADDRESSES
    # the module's settings are baked into its graphs: changing one re-captures
    for crit in (eager, graphed):
        crit.temperature, crit.contrast_mode = 0.2, 'one'
    for step in range(3):
        a = bufs[step % 2].detach().clone().requires_grad_(True)
        b = bufs[step % 2].detach().requires_grad_(True)
        la, lb = eager(a), graphed(b)
        la.backward()
        lb.backward()
        assert float(la) == float(lb)
        assert torch.equal(a.grad, b.grad)
