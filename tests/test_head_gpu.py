"""Projection-head tail: `ProjectionTail` against the expression it replaces,
`F.normalize(F.linear(h, W, b), dim=1)` (/root/reference/networks/resnet_big.py:168-172,177-181) evaluated by torch
in fp64 on the same bf16-rounded operands, forward and backward, and chained into the loss."""
import pytest
import torch
import torch.nn.functional as F

from simclr_pytorch_distributed_b200 import ProjectionTail, SupConLoss

pytestmark = pytest.mark.gpu


def make(M, K, N, dtype, seed=0, bias=True):
    g = torch.Generator().manual_seed(seed)
    h = torch.relu(torch.randn(M, K, generator=g)).to(dtype).cuda()
    lin = torch.nn.Linear(K, N, bias=bias)
    with torch.no_grad():
        lin.weight.copy_(torch.randn(N, K, generator=g) / K ** 0.5)
        if bias:
            lin.bias.copy_(torch.randn(N, generator=g) * 0.1)
    return h, lin.cuda()


def reference(h, lin, n_views, normalize=True):
    """fp64 on the operands the kernel sees (bf16-rounded h and W, fp32 bias)."""
    h64 = h.detach().to(torch.bfloat16).double().requires_grad_(True)
    w64 = lin.weight.detach().to(torch.bfloat16).double().requires_grad_(True)
    b64 = lin.bias.detach().double().requires_grad_(True) if lin.bias is not None else None
    y = F.linear(h64, w64, b64)
    if normalize:
        y = F.normalize(y, dim=1)
    if n_views > 1:
        y = y.view(n_views, -1, y.shape[1]).transpose(0, 1)          # main_supcon.py:285-286
    return y, h64, w64, b64


def rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


@pytest.mark.parametrize('M,K,N,n_views', [(1024, 2048, 128, 2), (1000, 2048, 128, 1), (256, 200, 64, 2),
                                            (128, 64, 128, 1), (77 * 3, 512, 32, 3), (4096, 2048, 128, 2)])
@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float32])
def test_forward_and_backward_match_torch(M, K, N, n_views, dtype):
    h, lin = make(M, K, N, dtype, seed=M + K)
    tail = ProjectionTail.from_linear(lin, n_views=n_views)
    hq = h.clone().requires_grad_(True)
    y = tail(hq)
    y_ref, h64, w64, b64 = reference(h, lin, n_views)
    assert y.shape == y_ref.shape and y.dtype == dtype
    if dtype == torch.float32:
        assert float((y.double() - y_ref).abs().max()) <= 2e-6            # fp32 accumulation order only
    else:
        assert float((y.double() - y_ref).abs().max()) <= 2.0 ** -8       # one bf16 ulp of a value below 1
    # unit rows
    assert float((y.float().reshape(-1, N).norm(dim=1) - 1).abs().max()) <= (1e-5 if dtype == torch.float32 else 8e-3)
    g = torch.randn(y.shape, generator=torch.Generator().manual_seed(5)).cuda()
    y.backward(g.to(dtype))
    y_ref.backward(g.to(dtype).double())
    tol = 1e-5 if dtype == torch.float32 else 1e-2       # bf16: the gradient GEMMs run in bf16 (what autocast does)
    assert rel(hq.grad, h64.grad) <= tol
    assert rel(lin.weight.grad, w64.grad) <= tol
    assert rel(lin.bias.grad, b64.grad) <= tol
    lin.weight.grad = None
    lin.bias.grad = None


def test_plain_linear_and_no_bias():
    h, lin = make(300, 256, 64, torch.float32, seed=3, bias=False)
    tail = ProjectionTail.from_linear(lin, normalize=False)
    y = tail(h)
    y_ref, _, _, _ = reference(h, lin, 1, normalize=False)
    assert float((y.double() - y_ref).abs().max()) <= 2e-5


def test_state_dict_is_a_linear_state_dict():
    lin = torch.nn.Linear(64, 32)
    tail = ProjectionTail(64, 32)
    tail.load_state_dict(lin.state_dict())
    lin.load_state_dict(tail.state_dict())
    assert list(tail.state_dict().keys()) == ['weight', 'bias']


def test_argument_errors():
    with pytest.raises(ValueError, match='out_features'):
        ProjectionTail(64, 256)
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        ProjectionTail(64, 32)(torch.zeros(4, 64))
    with pytest.raises(ValueError, match='multiple of n_views'):
        ProjectionTail(64, 32, n_views=2).cuda()(torch.zeros(3, 64, device='cuda'))


def test_chained_into_the_loss_matches_the_reference_chain():
    """head tail -> loss (bf16 rows used in place) against normalize(linear) -> stack -> dense loss in fp64."""
    from oracle import dense_port
    B, K, N = 512, 2048, 128
    h, lin = make(2 * B, K, N, torch.bfloat16, seed=11)
    tail = ProjectionTail.from_linear(lin, n_views=2)
    crit = SupConLoss(temperature=0.5)
    hq = h.clone().requires_grad_(True)
    z = tail(hq)
    assert z.is_contiguous() and z.shape == (B, 2, N)
    loss = crit(z)
    loss.backward()
    y_ref, h64, w64, b64 = reference(h, lin, 2)
    loss_ref = dense_port.dense_supcon_loss(y_ref, temperature=0.5, base_temperature=0.07)
    loss_ref.backward()
    assert abs(float(loss) - float(loss_ref)) <= 2e-3 * abs(float(loss_ref)) + 2.0 ** -8 * abs(float(loss_ref))
    assert rel(hq.grad, h64.grad) <= 2e-2
    assert rel(lin.weight.grad, w64.grad) <= 2e-2
