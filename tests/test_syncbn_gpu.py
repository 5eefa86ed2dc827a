"""Parity of the CUDA SyncBatchNorm kernels (through the module and through the raw C ABI) with the
oracle, the golden vectors, and torch's own batch-norm on the same GPU.  north_star tolerance:
outputs and gradients within 1e-5 (fp32 quantities); bf16 tensors are compared after the same
rounding, i.e. to within one bf16 ulp."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from conftest import golden
from oracle import syncbn_forward, syncbn_backward
from simclr_pytorch_distributed_b200 import SyncBatchNorm, _lib
from simclr_pytorch_distributed_b200.sync_batchnorm import _Kernels, _layout

pytestmark = pytest.mark.gpu
DEV = 'cuda'


def ours_fwd_bwd(x, dy, bn):
    xd = x.clone().requires_grad_(True)
    y = bn(xd)
    y.backward(dy)
    torch.cuda.synchronize()
    return y.detach(), xd.grad, bn.weight.grad.clone() if bn.weight is not None else None, \
        bn.bias.grad.clone() if bn.bias is not None else None


def oracle_fwd_bwd(x, dy, bn_state, eps=1e-5, momentum=0.1):
    w, b, rm, rv = bn_state
    xs = [x.detach().double().cpu().numpy()]
    ys, mean, invstd, n, rm1, rv1 = syncbn_forward(xs, w, b, eps, rm, rv, momentum)
    dxs, gws, gbs, _, _ = syncbn_backward(xs, [dy.double().cpu().numpy()], mean, invstd, w)
    return ys[0], dxs[0], gws[0], gbs[0], rm1, rv1, mean, invstd


def make_bn(C, **kw):
    bn = SyncBatchNorm(C, **kw).to(DEV)
    with torch.no_grad():
        if bn.weight is not None:
            bn.weight.uniform_(0.5, 1.5)
            bn.bias.normal_(0, 0.1)
        if bn.running_mean is not None:
            bn.running_mean.normal_(0, 0.3)
            bn.running_var.uniform_(0.5, 2.0)
    return bn


def state(bn):
    f = lambda t: None if t is None else t.detach().double().cpu().numpy().copy()   # noqa: E731
    return f(bn.weight), f(bn.bias), f(bn.running_mean), f(bn.running_var)


SHAPES = [((256, 256, 8, 8), True), ((32, 64, 32, 32), True), ((16, 2048, 4, 4), True),
          ((8, 24, 5, 7), True), ((7, 13, 3, 3), True), ((64, 128), False), ((4, 6, 3, 3), False),
          ((16, 64, 16, 16), False), ((5, 10, 7, 9), False), ((3, 4, 2, 3, 5), False)]


@pytest.mark.parametrize('shape,cl', SHAPES)
def test_fp32_matches_oracle_1e5(shape, cl):
    torch.manual_seed(sum(shape))
    C = shape[1]
    mu = (torch.rand(C) * 4 - 2).view(1, C, *([1] * (len(shape) - 2)))
    sg = (torch.rand(C) * 1.5 + 0.5).view(1, C, *([1] * (len(shape) - 2)))
    x = (torch.randn(shape) * sg + mu).to(DEV)
    dy = torch.randn(shape).to(DEV)
    if cl:
        x = x.contiguous(memory_format=torch.channels_last)
        dy = dy.contiguous(memory_format=torch.channels_last)
    bn = make_bn(C)
    st = state(bn)
    y, dx, gw, gb = ours_fwd_bwd(x, dy, bn)
    ry, rdx, rgw, rgb, rm1, rv1, _, _ = oracle_fwd_bwd(x, dy, st)
    assert y.shape == x.shape and y.stride() == x.stride()
    assert np.abs(y.cpu().numpy() - ry).max() <= 1e-5 * max(1.0, np.abs(ry).max())
    assert np.abs(dx.cpu().numpy() - rdx).max() <= 1e-5 * max(1.0, np.abs(rdx).max())
    n = x.numel() // C
    assert np.abs(gw.cpu().numpy() - rgw).max() <= 1e-5 * max(1.0, np.abs(rgw).max()) * max(1, n ** 0.5 / 8)
    assert np.abs(gb.cpu().numpy() - rgb).max() <= 1e-5 * max(1.0, np.abs(rgb).max()) * max(1, n ** 0.5 / 8)
    assert np.abs(bn.running_mean.cpu().numpy() - rm1).max() <= 1e-5
    assert np.abs(bn.running_var.cpu().numpy() - rv1).max() <= 1e-5 * max(1.0, np.abs(rv1).max())
    assert int(bn.num_batches_tracked) == 1


@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float16])
@pytest.mark.parametrize('shape,cl', SHAPES[:5] + SHAPES[6:8])
def test_half_precision_io_within_one_ulp(shape, cl, dtype):
    torch.manual_seed(sum(shape) + 1)
    C = shape[1]
    x = (torch.randn(shape) * 1.3 + 0.4).to(DEV, dtype)
    dy = torch.randn(shape).to(DEV, dtype)
    if cl:
        x = x.contiguous(memory_format=torch.channels_last)
        dy = dy.contiguous(memory_format=torch.channels_last)
    bn = make_bn(C)
    st = state(bn)
    y, dx, gw, gb = ours_fwd_bwd(x, dy, bn)
    assert y.dtype == dtype and dx.dtype == dtype and gw.dtype == torch.float32
    ry, rdx, rgw, rgb, rm1, rv1, mean, invstd = oracle_fwd_bwd(x, dy, st)
    ulp = 2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -11
    assert np.all(np.abs(y.float().cpu().numpy() - ry) <= ulp * np.abs(ry) + 1e-5)
    assert np.all(np.abs(dx.float().cpu().numpy() - rdx) <= ulp * np.abs(rdx) + 1e-5)
    # fp32 per-channel quantities keep the 1e-5 bar
    assert np.abs(bn.running_mean.cpu().numpy() - rm1).max() <= 1e-5
    assert np.abs(bn.running_var.cpu().numpy() - rv1).max() <= 1e-5 * max(1.0, np.abs(rv1).max())
    n = x.numel() // C
    assert np.abs(gb.cpu().numpy() - rgb).max() <= 1e-5 * max(1.0, np.abs(rgb).max()) * max(1, n ** 0.5 / 8)


def test_matches_torch_batch_norm_on_gpu():
    """torch's own kernels on the same device (F.batch_norm == nn.SyncBatchNorm at world size 1,
    torch/nn/modules/batchnorm.py:818-828)."""
    torch.manual_seed(11)
    for shape, cl in SHAPES[:4] + SHAPES[5:8]:
        C = shape[1]
        x = (torch.randn(shape, device=DEV) * 2 + 1)
        dy = torch.randn(shape, device=DEV)
        if cl:
            x = x.contiguous(memory_format=torch.channels_last)
        bn = make_bn(C)
        ref = torch.nn.BatchNorm2d(C).to(DEV) if len(shape) == 4 else torch.nn.BatchNorm1d(C).to(DEV)
        ref.load_state_dict(bn.state_dict())
        y, dx, gw, gb = ours_fwd_bwd(x, dy, bn)
        xr = x.clone().requires_grad_(True)
        yr = ref(xr)
        yr.backward(dy)
        assert torch.allclose(y, yr, atol=2e-5, rtol=1e-5)
        assert torch.allclose(dx, xr.grad, atol=2e-5, rtol=1e-5)
        assert torch.allclose(gw, ref.weight.grad, atol=1e-3, rtol=1e-4)
        assert torch.allclose(bn.running_var, ref.running_var, atol=1e-5, rtol=1e-5)


def test_matches_aten_sync_ops_on_gpu():
    """The very ops stock SyncBatchNorm calls (_functions.py:39,122,144,170), NHWC bf16 cfg 2."""
    torch.manual_seed(12)
    x = (torch.randn(256, 256, 8, 8, device=DEV) * 1.5 + 0.5).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    dy = torch.randn(256, 256, 8, 8, device=DEV).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    bn = make_bn(256)
    w, b = bn.weight.detach(), bn.bias.detach()
    mean, invstd = torch.batch_norm_stats(x, 1e-5)
    y_ref = torch.batch_norm_elemt(x, w, b, mean, invstd, 1e-5)
    sum_dy, sum_dy_xmu, gw_ref, gb_ref = torch.batch_norm_backward_reduce(dy, x, mean, invstd, w, True, True, True)
    count = torch.full((1,), x.numel() // 256, dtype=torch.int32, device=DEV)
    dx_ref = torch.batch_norm_backward_elemt(dy, x, mean, invstd, w, sum_dy, sum_dy_xmu, count)
    y, dx, gw, gb = ours_fwd_bwd(x, dy, bn)
    ulp = 2.0 ** -7      # both sides are rounded to bf16 independently: they may land one ulp apart
    assert torch.all((y.float() - y_ref.float()).abs() <= ulp * y_ref.float().abs() + 1e-5)
    assert torch.all((dx.float() - dx_ref.float()).abs() <= ulp * dx_ref.float().abs() + 1e-5)
    assert torch.allclose(gw, gw_ref, rtol=1e-4, atol=1e-3)
    assert torch.allclose(gb, gb_ref, rtol=1e-4, atol=1e-3)


@pytest.mark.parametrize('name', ['bn_w1', 'bn_w2', 'bn_w3_ragged', 'bn_w2_f32'])
@pytest.mark.parametrize('cl', [False, True])
def test_golden_multi_rank_arithmetic_through_c_abi(name, cl):
    """The W-rank exchange emulated on one GPU: each rank's kernels run one after the other, the
    messages are concatenated as the all-gather would (stats_all [W, 2C+1]) and the 2C sums added
    as the all-reduce would.  Ragged per-rank counts included (Chan merge with counts)."""
    g = golden(name)
    W = len(g['counts'])
    K = _Kernels
    f32 = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(DEV, torch.float32)   # noqa: E731
    xs = [f32(g[f'x{r}']) for r in range(W)]
    dys = [f32(g[f'dy{r}']) for r in range(W)]
    if cl:
        xs = [t.contiguous(memory_format=torch.channels_last) for t in xs]
        dys = [t.contiguous(memory_format=torch.channels_last) for t in dys]
    C = xs[0].shape[1]
    w, b = f32(g['weight']), f32(g['bias'])
    msgs = []
    for x in xs:
        x_, geom = _layout(x)
        st = torch.empty(2 * C + 1, device=DEV)
        K.stats(x_, geom, st)
        msgs.append(st)
    stats_all = torch.cat(msgs)
    tol = 2e-5
    saved_all, sums_tot = [], torch.zeros(2 * C, device=DEV)
    gws, gbs = [], []
    for r, x in enumerate(xs):
        x_, geom = _layout(x)
        rm, rv = f32(g['running_mean0']), f32(g['running_var0'])
        y = torch.empty_like(x_)
        saved = torch.empty(2 * C + 1, device=DEV)
        K.apply(x_, y, geom, stats_all, W, w, b, rm, rv, float(g['momentum']), float(g['eps']), saved, 0)
        assert np.abs(y.cpu().numpy() - g[f'y{r}']).max() <= tol * max(1, np.abs(g[f'y{r}']).max())
        assert np.abs(rm.cpu().numpy() - g['running_mean1']).max() <= tol
        assert np.abs(rv.cpu().numpy() - g['running_var1']).max() <= tol
        saved_all.append(saved)
        out = torch.empty(4 * C, device=DEV)
        K.backward_reduce(dys[r], x_, geom, saved, w, b, out[:2 * C], out[2 * C:3 * C], out[3 * C:], 0)
        sums_tot += out[:2 * C]
        gws.append(out[2 * C:3 * C].clone())
        gbs.append(out[3 * C:].clone())
    for r, x in enumerate(xs):
        x_, geom = _layout(x)
        dx = torch.empty_like(x_)
        K.backward_apply(dys[r], x_, dx, geom, saved_all[r], w, b, sums_tot, 0)
        assert np.abs(dx.cpu().numpy() - g[f'dx{r}']).max() <= tol * max(1, np.abs(g[f'dx{r}']).max())
    assert np.abs(sum(gws).cpu().numpy() - g['grad_weight']).max() <= 1e-4 * max(1, np.abs(g['grad_weight']).max())
    assert np.abs(sum(gbs).cpu().numpy() - g['grad_bias']).max() <= 1e-4 * max(1, np.abs(g['grad_bias']).max())


def test_module_variants():
    torch.manual_seed(13)
    x = torch.randn(16, 32, 6, 6, device=DEV) * 2 + 1
    dy = torch.randn_like(x)
    # affine=False
    bn = make_bn(32, affine=False)
    xd = x.clone().requires_grad_(True)
    y = bn(xd)
    y.backward(dy)
    yr = F.batch_norm(x, None, None, None, None, True, 0.1, 1e-5)
    assert torch.allclose(y, yr, atol=2e-5)
    # momentum=None: cumulative moving average (batchnorm.py:760-767)
    bn = make_bn(32, momentum=None)
    ref = torch.nn.BatchNorm2d(32, momentum=None).to(DEV)
    ref.load_state_dict(bn.state_dict())
    for _ in range(3):
        bn(x)
        ref(x)
    assert torch.allclose(bn.running_mean, ref.running_mean, atol=1e-5)
    assert torch.allclose(bn.running_var, ref.running_var, atol=1e-5)
    assert int(bn.num_batches_tracked) == 3
    # track_running_stats=False: batch stats in eval too
    bn = make_bn(32, track_running_stats=False).eval()
    assert torch.allclose(bn(x), F.batch_norm(x, None, None, bn.weight, bn.bias, True, 0.0, 1e-5), atol=2e-5)
    # eval mode uses running stats and leaves them alone; gradients still flow
    bn = make_bn(32).eval()
    rm0 = bn.running_mean.clone()
    xd = x.clone().requires_grad_(True)
    y = bn(xd)
    y.backward(dy)
    xr = x.clone().requires_grad_(True)
    yr = F.batch_norm(xr, bn.running_mean, bn.running_var, bn.weight, bn.bias, False, 0.0, 1e-5)
    yr.backward(dy)
    assert torch.allclose(y, yr, atol=2e-5) and torch.equal(bn.running_mean, rm0)
    assert torch.allclose(xd.grad, xr.grad, atol=2e-5)
    # one value per channel in training mode
    with pytest.raises(ValueError, match='Expected more than 1 value per channel'):
        make_bn(4)(torch.randn(1, 4, device=DEV))
    # repeated calls reuse the self-cleaning scratch
    bn = make_bn(32)
    a = bn(x)
    bn2 = make_bn(32)
    bn2.load_state_dict({k: v for k, v in bn.state_dict().items()})
    c = bn(x)
    # the cross-CTA sums use float atomics: the order, hence the last bit, may differ between calls
    assert torch.allclose(a, c, atol=1e-5, rtol=0)


def test_fused_relu():
    torch.manual_seed(14)
    for cl in (True, False):
        x = torch.randn(8, 64, 10, 10, device=DEV)
        dy = torch.randn_like(x)
        if cl:
            x = x.contiguous(memory_format=torch.channels_last)
        bn = make_bn(64, fuse_relu=True)
        ref = torch.nn.BatchNorm2d(64).to(DEV)
        ref.load_state_dict(bn.state_dict())
        y, dx, gw, gb = ours_fwd_bwd(x, dy, bn)
        xr = x.clone().requires_grad_(True)
        yr = torch.relu(ref(xr))
        yr.backward(dy)
        assert torch.allclose(y, yr, atol=2e-5)
        assert torch.allclose(dx, xr.grad, atol=5e-5)
        assert torch.allclose(gw, ref.weight.grad, atol=1e-3, rtol=1e-4)
        assert torch.allclose(gb, ref.bias.grad, atol=1e-3, rtol=1e-4)


def test_convert_and_train_small_cnn_like_reference_swap():
    """main_supcon.py:222-224: convert_sync_batchnorm(model) then train."""
    torch.manual_seed(15)
    def net():
        return torch.nn.Sequential(torch.nn.Conv2d(3, 16, 3, padding=1, bias=False), torch.nn.BatchNorm2d(16),
                                   torch.nn.ReLU(), torch.nn.Conv2d(16, 32, 3, padding=1, bias=False),
                                   torch.nn.BatchNorm2d(32), torch.nn.ReLU(), torch.nn.AdaptiveAvgPool2d(1),
                                   torch.nn.Flatten(), torch.nn.Linear(32, 8))
    a = net().to(DEV)
    b = net().to(DEV)
    b.load_state_dict(a.state_dict())
    b = SyncBatchNorm.convert_sync_batchnorm(b)
    x = torch.randn(32, 3, 16, 16, device=DEV)
    la = a(x).square().mean()
    lb = b(x).square().mean()
    la.backward()
    lb.backward()
    assert torch.allclose(la, lb, rtol=1e-5)
    for pa, pb in zip(a.parameters(), b.parameters()):
        assert torch.allclose(pa.grad, pb.grad, atol=1e-5, rtol=1e-4)
    assert list(a.state_dict().keys()) == list(b.state_dict().keys())


def test_large_layer_shapes_property():
    """ResNet-50 layer shapes at N=512 are too big for a CPU oracle in a unit
    test: check size-independent properties instead -- per-channel mean 0 / variance 1 of the
    output, sum of dx per channel = 0 and sum of dx*xhat = 0 (the BN backward projections)."""
    torch.manual_seed(16)
    for shape in [(512, 64, 32, 32), (512, 1024, 8, 8)]:
        C = shape[1]
        x = (torch.randn(shape, device=DEV, dtype=torch.float32) * 1.7 + 0.3).contiguous(memory_format=torch.channels_last)
        dy = torch.randn(shape, device=DEV).contiguous(memory_format=torch.channels_last)
        bn = SyncBatchNorm(C, affine=False).to(DEV)
        xd = x.clone().requires_grad_(True)
        y = bn(xd)
        y.backward(dy)
        m = y.double().mean(dim=(0, 2, 3))
        v = y.double().var(dim=(0, 2, 3), unbiased=False)
        assert m.abs().max() < 1e-5 and (v - 1).abs().max() < 1e-4
        assert xd.grad.double().mean(dim=(0, 2, 3)).abs().max() < 1e-6
        assert (xd.grad.double() * y.double()).mean(dim=(0, 2, 3)).abs().max() < 1e-5


# ----------------------------------------------------------------------------- residual add + ReLU (resnet_big.py:27-30,60-63)
@pytest.mark.parametrize('shape,cl,dtype', [((32, 64, 16, 16), True, torch.float32), ((256, 256, 8, 8), True, torch.bfloat16),
                                            ((512, 64, 32, 32), True, torch.bfloat16), ((16, 24, 5, 7), False, torch.float32),
                                            ((64, 128), False, torch.float32),
                                            ((640, 128, 8, 8), True, torch.bfloat16),      # forward: shared-memory-resident pass
                                            ((320, 128, 8, 8), True, torch.bfloat16)])     # both directions resident
@pytest.mark.parametrize('relu', [True, False])
def test_residual_add_relu_matches_unfused_expression(shape, cl, dtype, relu):
    """y = relu(bn(x) + z) in one pass vs torch autograd on the unfused expression (fp64 reference)."""
    torch.manual_seed(sum(shape) + int(relu))
    C = shape[1]
    x = (torch.randn(shape, device=DEV) * 1.5 + 0.3).to(dtype)
    z = torch.randn(shape, device=DEV).to(dtype)
    dy = torch.randn(shape, device=DEV).to(dtype)
    if cl:
        x, z, dy = (t.contiguous(memory_format=torch.channels_last) for t in (x, z, dy))
    bn = make_bn(C, fuse_relu=relu)
    ref = (torch.nn.BatchNorm2d(C) if len(shape) == 4 else torch.nn.BatchNorm1d(C)).to(DEV).double()
    ref.load_state_dict(bn.state_dict())
    xd, zd = x.clone().requires_grad_(True), z.clone().requires_grad_(True)
    y = bn(xd, zd)
    y.backward(dy)
    xr, zr = x.double().requires_grad_(True), z.double().requires_grad_(True)
    yr = ref(xr) + zr
    if relu:
        # the gate of a pre-activation within fp32 rounding of zero is a coin toss: take OUR gate for the reference
        # (it is the sign of the value we computed) and check that the two only differ there
        pre = yr.detach()
        ours_on = y.detach() > 0
        assert torch.all((ours_on == (pre > 0)) | (pre.abs() < 1e-5))
        yr = yr * ours_on
    yr.backward(dy.double())
    if dtype == torch.float32:
        assert torch.allclose(y.double(), yr, atol=2e-5)
        assert torch.allclose(xd.grad.double(), xr.grad, atol=5e-5)
        assert torch.allclose(zd.grad.double(), zr.grad, atol=1e-6)
    else:
        ulp = 2.0 ** -7
        assert torch.all((y.double() - yr).abs() <= ulp * yr.abs() + 2e-2)
        near0 = yr.abs() < 2e-2
        assert torch.all(((xd.grad.double() - xr.grad).abs() <= ulp * xr.grad.abs() + 2e-2) | near0)
        assert torch.all((zd.grad.double() == zr.grad) | near0)
    assert torch.allclose(bn.weight.grad.double(), ref.weight.grad, rtol=2e-3 if dtype != torch.float32 else 1e-4,
                          atol=5e-2 if dtype != torch.float32 else 1e-3)


@pytest.mark.parametrize('shape,dtype', [((256, 256, 8, 8), torch.bfloat16), ((64, 128, 8, 8), torch.float32),
                                         ((512, 512, 4, 4), torch.bfloat16), ((128, 256, 16, 16), torch.bfloat16),
                                         ((384, 256, 8, 8), torch.bfloat16), ((512, 256, 8, 8), torch.bfloat16),
                                         ((200, 96, 8, 8), torch.float32),
                                         # ragged row counts through the shared-memory-parked backward (32-byte slabs x clusters of 6)
                                         ((250, 256, 8, 8), torch.bfloat16), ((255, 256, 7, 9), torch.float16)])
def test_cluster_kernels_match_multi_kernel_path(shape, dtype, monkeypatch):
    """Single-launch kernels (register-resident cluster passes; for the 12-16 MB shapes the shared-memory-resident
    passes behind the cooperative grid barrier) vs stats -> apply / reduce -> apply on the same tensors."""
    torch.manual_seed(5)
    C = shape[1]
    x = (torch.randn(shape, device=DEV) * 1.7 + 0.4).to(dtype).contiguous(memory_format=torch.channels_last)
    dy = torch.randn(shape, device=DEV).to(dtype).contiguous(memory_format=torch.channels_last)
    # Keep every pre-activation away from the ReLU threshold.  The passes that sum through float atomics are not
    # bitwise reproducible, so an element within rounding of zero may be gated differently from run to run, and ONE
    # flipped element moves sum(dy), dgamma and dbeta of its channel by O(1) (seen on channel 52 of [384,256,8,8]).
    ax = [0, 2, 3]
    xd64 = x.double()
    pre = (xd64 - xd64.mean(ax, keepdim=True)) / torch.sqrt(xd64.var(ax, unbiased=False, keepdim=True) + 1e-5) * 1.25 + 0.1
    x = torch.where(pre.abs() < 4e-3, x + 0.25, x).contiguous(memory_format=torch.channels_last)
    outs = []
    for fused in ('1', '0'):
        monkeypatch.setenv('SCB_BN_FUSED', fused)
        bn = make_bn(C, fuse_relu=True)
        with torch.no_grad():
            bn.running_mean.zero_(); bn.running_var.fill_(1.0)
            bn.weight.fill_(1.25); bn.bias.fill_(0.1)
        y, dx, gw, gb = ours_fwd_bwd(x, dy, bn)
        outs.append((y.float(), dx.float(), gw, gb, bn.running_mean.clone(), bn.running_var.clone()))
    tol = 2.0 ** -7 if dtype == torch.bfloat16 else (2.0 ** -10 if dtype == torch.float16 else 1e-5)
    # the ReLU gate may flip for outputs within rounding of zero: compare away from it
    far = outs[1][0].abs() > 1e-2
    assert torch.all(((outs[0][0] - outs[1][0]).abs() <= tol * outs[1][0].abs() + 1e-5) | ~far)
    assert torch.all(((outs[0][1] - outs[1][1]).abs() <= tol * outs[1][1].abs() + 1e-4) | ~far)
    for a, b in zip(outs[0][2:], outs[1][2:]):
        assert torch.allclose(a, b, rtol=1e-4, atol=1e-3 * max(1.0, float(b.abs().max())))


def test_fused_supported_reports_capacity():
    """[256,256,8,8] bf16 fits forward and backward; a 268 MB layer does not."""
    x = torch.empty(256, 256, 8, 8, device=DEV, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    x_, geom = _layout(x)
    assert _Kernels.fused_supported(x_, geom, 0) and _Kernels.fused_supported(x_, geom, 1)
    assert not _Kernels.fused_supported(x_, (1, 512 * 32 * 32, 256, 1), 1)


@pytest.mark.parametrize('shape,cl,dtype', [
    ((256, 256, 8, 8), True, torch.bfloat16),      # register-resident cluster pass
    ((512, 256, 8, 8), True, torch.bfloat16),      # shared-memory-resident pass (cooperative grid barrier)
    ((512, 128, 16, 16), True, torch.bfloat16),    # ring passes
    ((8, 12, 5, 5), True, torch.float32),          # tiled passes (rows are not a whole number of 16-byte vectors)
    ((8, 16, 6, 6), False, torch.float32),         # NCHW
])
def test_step_counter_is_kept_by_the_kernels(shape, cl, dtype):
    """num_batches_tracked (batchnorm.py:763-764) is incremented by the normalising kernel itself: once per training
    forward on every kernel family, never in eval mode."""
    bn = make_bn(shape[1])
    x = torch.randn(shape, device=DEV).to(dtype)
    if cl:
        x = x.contiguous(memory_format=torch.channels_last)
    for i in range(3):
        bn(x)
        assert int(bn.num_batches_tracked) == i + 1
    bn.eval()
    bn(x)
    assert int(bn.num_batches_tracked) == 3
    bn.train()
    bn(x, torch.zeros_like(x))                      # the residual-add variants count too
    assert int(bn.num_batches_tracked) == 4


@pytest.mark.parametrize('shape,cl', [((256, 256, 8, 8), True), ((512, 256, 8, 8), True), ((512, 128, 16, 16), True),
                                      ((8, 12, 5, 5), True), ((16, 64, 16, 16), False)])
@pytest.mark.parametrize('pdtype', [torch.bfloat16, torch.float16])
def test_low_precision_parameters_are_taken_natively(shape, cl, pdtype):
    """A module converted with .to(bf16 / fp16) -- weight, bias and running statistics in that dtype, as
    `model.to(torch.bfloat16)` leaves them -- against the same module with fp32 copies of the same parameter values:
    same y and dx, running statistics and parameter gradients equal up to the rounding of their own dtype, and no
    change of dtype anywhere (torch's batch norm takes mixed dtypes the same way)."""
    torch.manual_seed(11)
    C = shape[1]
    x = (torch.randn(shape, device=DEV) * 1.3 + 0.2).to(torch.bfloat16)
    dy = torch.randn(shape, device=DEV).to(torch.bfloat16)
    if cl:
        x = x.contiguous(memory_format=torch.channels_last)
        dy = dy.contiguous(memory_format=torch.channels_last)
    low = make_bn(C).to(pdtype)
    ref = make_bn(C)
    ref.load_state_dict({k: v.float() if v.is_floating_point() else v for k, v in low.state_dict().items()})
    outs = []
    for bn in (low, ref):
        xd = x.clone().requires_grad_(True)
        y = bn(xd)
        y.backward(dy)
        outs.append((y.detach().float(), xd.grad.float(), bn.weight.grad, bn.bias.grad, bn.running_mean, bn.running_var))
    assert low.weight.dtype == pdtype and low.running_mean.dtype == pdtype
    assert outs[0][2].dtype == pdtype and outs[0][3].dtype == pdtype
    ulp = 2.0 ** -7 if pdtype == torch.bfloat16 else 2.0 ** -10
    assert float((outs[0][0] - outs[1][0]).abs().max()) <= 2.0 ** -7 * float(outs[1][0].abs().max())
    assert float((outs[0][1] - outs[1][1]).abs().max()) <= 2.0 ** -7 * float(outs[1][1].abs().max()) + 1e-4
    for a, b in zip(outs[0][2:], outs[1][2:]):
        assert torch.all((a.float() - b.float()).abs() <= ulp * b.float().abs() + 1e-3 * float(b.float().abs().max()) + 1e-6)
    assert int(low.num_batches_tracked) == 1
