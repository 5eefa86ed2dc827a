"""Host-side behaviour of the drop-in modules that needs no GPU: argument checks mirror the
reference's (`/root/reference/losses.py:33-61`, torch/nn/modules/batchnorm.py:735-743), CPU tensors
are refused instead of silently taking another path, module state matches torch's."""
import pytest
import torch

from simclr_pytorch_distributed_b200 import SupConLoss, SyncBatchNorm
from simclr_pytorch_distributed_b200 import losses as L
from simclr_pytorch_distributed_b200 import _lib


def test_constructor_defaults_match_reference():
    c = SupConLoss()
    assert (c.temperature, c.contrast_mode, c.base_temperature) == (0.07, 'all', 0.07)
    c = SupConLoss(temperature=0.5)           # main_supcon.py:214: base_temperature stays 0.07
    assert c.base_temperature == 0.07
    assert list(c.parameters()) == [] and list(c.buffers()) == []


def test_value_errors_in_reference_order():
    c = SupConLoss()
    with pytest.raises(ValueError, match='at least 3 dimensions are required'):
        c(torch.zeros(4, 8))
    f = torch.zeros(4, 2, 8)
    with pytest.raises(ValueError, match='Cannot define both `labels` and `mask`'):
        c(f, labels=torch.arange(4), mask=torch.eye(4))
    with pytest.raises(ValueError, match='Num of labels does not match num of features'):
        c(f, labels=torch.arange(3))
    with pytest.raises(ValueError, match='Unknown mode: two'):
        SupConLoss(contrast_mode='two')(f)


def test_cpu_tensor_is_refused_loudly():
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        SupConLoss()(torch.zeros(4, 2, 8))
    bn = SyncBatchNorm(8)
    with pytest.raises(ValueError, match='expected input tensor to be on GPU'):
        bn(torch.zeros(4, 8, 2, 2))


def test_engine_selection():
    assert L.select_engine('auto', torch.float32, 128, _lib.MASK_SIMCLR, 512) == _lib.ENGINE_BF16X3
    assert L.select_engine('auto', torch.float32, 128, _lib.MASK_SIMCLR, 512, allow_x3=False) == _lib.ENGINE_FP32
    assert L.select_engine('auto', torch.bfloat16, 128, _lib.MASK_LABELS, 512) == _lib.ENGINE_BF16
    assert L.select_engine('auto', torch.float32, 128, _lib.MASK_EXPLICIT, 512) == _lib.ENGINE_FP32
    assert L.select_engine('auto', torch.float32, 4096, _lib.MASK_SIMCLR, 512) == _lib.ENGINE_FP32
    with pytest.raises(ValueError, match='Unknown engine'):
        L.select_engine('fp64', torch.float32, 128, _lib.MASK_SIMCLR, 512)
    with pytest.raises(ValueError, match='does not take this problem'):
        L.select_engine('bf16', torch.float32, 128, _lib.MASK_EXPLICIT, 512)


def test_loss_plan_row_bookkeeping():
    p = L.LossPlan(n_local=8, n_views=2, dim=24, n_seg=4, seg_index=2, mode=_lib.MODE_ONE,
                   engine=_lib.ENGINE_BF16)
    assert p.rows_local == 16 and p.n_k == 64 and p.row_begin == 32
    assert p.n_anchor_total == 32 and p.ldk == 64
    p = L.LossPlan(n_local=8, n_views=3, dim=24)
    assert p.n_anchor_total == 24 and p.ldk == 24 and p.row_begin == 0


def test_syncbn_input_checks_and_state():
    bn = SyncBatchNorm(8)
    with pytest.raises(ValueError, match='expected at least 2D input'):
        bn(torch.zeros(8))
    with pytest.raises(ValueError, match='channels should be non-zero'):
        bn(torch.zeros(4, 0, 2))
    ref = torch.nn.SyncBatchNorm(8)
    assert list(bn.state_dict().keys()) == list(ref.state_dict().keys())
    for k, v in ref.state_dict().items():
        assert bn.state_dict()[k].shape == v.shape and bn.state_dict()[k].dtype == v.dtype
    bn.load_state_dict(ref.state_dict())
    ref.load_state_dict(bn.state_dict())
    bn2 = SyncBatchNorm(8, affine=False, track_running_stats=False, momentum=None)
    assert bn2.weight is None and bn2.running_mean is None and bn2.momentum is None


def test_convert_sync_batchnorm_shares_parameters():
    net = torch.nn.Sequential(torch.nn.Conv2d(3, 8, 3), torch.nn.BatchNorm2d(8), torch.nn.ReLU(),
                              torch.nn.Sequential(torch.nn.Conv2d(8, 4, 1), torch.nn.BatchNorm2d(4)))
    w = net[1].weight
    rm = net[3][1].running_mean
    net[1].eval()
    out = SyncBatchNorm.convert_sync_batchnorm(net)
    assert isinstance(out[1], SyncBatchNorm) and isinstance(out[3][1], SyncBatchNorm)
    assert out[1].weight is w and out[3][1].running_mean is rm
    assert out[1].training is False and out[3][1].training is True
    assert list(out.state_dict().keys()) == list(net.state_dict().keys())
    # converting the stock torch.nn.SyncBatchNorm works too (it is a _BatchNorm)
    out2 = SyncBatchNorm.convert_sync_batchnorm(torch.nn.Sequential(torch.nn.SyncBatchNorm(4)))
    assert isinstance(out2[0], SyncBatchNorm)


def test_projection_tail_host_checks_and_state():
    from simclr_pytorch_distributed_b200 import ProjectionTail
    with pytest.raises(ValueError, match='out_features'):
        ProjectionTail(64, 256)
    with pytest.raises(ValueError, match='in_features'):
        ProjectionTail(60, 32)
    tail = ProjectionTail(64, 32, n_views=2)
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        tail(torch.zeros(4, 64))
    with pytest.raises(ValueError, match=r'expected input \[rows, 64\]'):
        tail(torch.zeros(4, 32))
    lin = torch.nn.Linear(64, 32)
    shared = ProjectionTail.from_linear(lin, n_views=2)
    assert shared.weight is lin.weight and shared.bias is lin.bias
    assert list(tail.state_dict().keys()) == list(lin.state_dict().keys())
    tail.load_state_dict(lin.state_dict())


def test_graph_mode_is_off_the_cpu_path():
    # graph=True changes nothing about argument checking: CPU tensors are still refused before any capture
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        SupConLoss(graph=True)(torch.zeros(4, 2, 8, requires_grad=True))
