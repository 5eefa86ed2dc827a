"""Single-GPU coverage of the peer-memory entry points and of the two row layouts of the loss.

On one GPU the "peers" are plain local buffers: the kernels, the flag protocol and the host plumbing are
the same as on an NVLink box (tests/test_multigpu.py covers real peers when 2+ GPUs are visible)."""
import ctypes

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import supcon_forward_backward
from simclr_pytorch_distributed_b200 import _lib
from simclr_pytorch_distributed_b200 import losses as L

pytestmark = pytest.mark.gpu


def _stream():
    return torch.cuda.current_stream().cuda_stream


def test_peer_bcast_copies_every_segment_to_every_destination():
    lib = _lib.load()
    src = torch.arange(4096, dtype=torch.float32, device='cuda')
    dsts = [torch.full((8192,), -1.0, device='cuda') for _ in range(3)]
    ptrs = (ctypes.c_void_p * 3)(*[d.data_ptr() for d in dsts])
    # 16-byte aligned segments (vector path) and a 4-byte granular one (scalar path) in separate calls
    seg = (ctypes.c_int64 * 6)(0, 1024 * 4, 512 * 4, 2048 * 4, 4096 * 4, 1024 * 4)
    _lib.check(lib.scb_peer_bcast(src.data_ptr(), seg, 2, ptrs, 3, _stream()), 'scb_peer_bcast')
    seg1 = (ctypes.c_int64 * 3)(4 * 3, 4 * 7001, 4 * 101)
    _lib.check(lib.scb_peer_bcast(src.data_ptr(), seg1, 1, ptrs, 3, _stream()), 'scb_peer_bcast')
    torch.cuda.synchronize()
    for d in dsts:
        assert torch.equal(d[1024:1536], src[0:512])
        assert torch.equal(d[4096:5120], src[2048:3072])
        assert torch.equal(d[7001:7102], src[3:104])
        assert float(d[0]) == -1.0 and float(d[1536]) == -1.0 and float(d[7000]) == -1.0 and float(d[7102]) == -1.0
    # misaligned arguments are refused, not rounded
    bad = (ctypes.c_int64 * 3)(2, 0, 16)
    assert lib.scb_peer_bcast(src.data_ptr(), bad, 1, ptrs, 3, _stream()) != 0
    assert b'multiples of 4' in lib.scb_last_error()


@pytest.mark.parametrize('engine', ['fp32', 'bf16'])
def test_fused_exchange_entry_points_with_local_peers(engine):
    """prep_peers -> wait -> forward_peers -> wait(+loss) with W = 1: same numbers as the plain calls,
    twice in a row (the counters advance, the flags follow)."""
    lib = _lib.load()
    dev = torch.device('cuda')
    B, V, D = 192, 2, 128
    torch.manual_seed(5)
    dtype = torch.float32 if engine == 'fp32' else torch.bfloat16
    feats = (torch.randn(B, V, D, device=dev) * 2 + 0.3).to(dtype)
    plan = L.LossPlan(B, V, D, engine=L._ENGINES[engine])
    plan.scalar_dtype = _lib.dtype_code(dtype)
    n_k, ldk = plan.n_k, plan.ldk
    k = torch.zeros((n_k, ldk), dtype=plan.k_torch, device=dev)
    st = torch.zeros(2 * n_k + 4, dtype=torch.float32, device=dev)
    flags = torch.zeros(32, dtype=torch.int32, device=dev)
    state = torch.zeros(4, dtype=torch.int32, device=dev)
    inv_norm = torch.empty(n_k, dtype=torch.float32, device=dev)
    possum = torch.empty(n_k, dtype=torch.float32, device=dev)
    loss2 = torch.empty(2, dtype=torch.float32, device=dev)
    chans = []
    for c, buf in ((0, k), (1, st)):
        ch = _lib.PeerChannel()
        ch.dst[0] = buf.data_ptr()
        ch.flag[0] = flags.data_ptr() + c * 64
        ch.state = state.data_ptr() + c * 8
        ch.n_dst = 1
        chans.append(ch)
    # reference: the plain entry points
    k_ref = torch.empty_like(k)
    inv_ref = torch.empty_like(inv_norm)
    L.launch_prep(lib, feats, plan, True, k_ref, None, inv_ref)
    prob_ref = L._problem(plan, k_ref, None, None, None, 0.5, 0.07)
    stats = torch.empty((3, n_k), dtype=torch.float32, device=dev)
    loss_ref = torch.empty(2, dtype=torch.float32, device=dev)
    L.launch_forward(lib, prob_ref, plan, stats[0], stats[1], stats[2], loss_ref, dev)
    for it in range(2):
        k.zero_()
        st.zero_()
        _lib.check(lib.scb_supcon_prep_peers(feats.data_ptr(), _lib.dtype_code(dtype), B, V, D, feats.stride(0),
                                             feats.stride(1), 1, 1e-12, ctypes.byref(chans[0]), 0, plan.k_dtype, ldk,
                                             inv_norm.data_ptr(), _stream()), 'scb_supcon_prep_peers')
        _lib.check(lib.scb_peer_wait(flags.data_ptr(), state.data_ptr(), 1, None, None, 0, _stream()), 'scb_peer_wait')
        prob = L._problem(plan, k, None, None, None, 0.5, 0.07)
        need = lib.scb_supcon_workspace_bytes(ctypes.byref(prob), n_k)
        ws = L._Workspace.get(dev, need)
        _lib.check(lib.scb_supcon_forward_peers(ctypes.byref(prob), 0, n_k, plan.n_anchor_total, possum.data_ptr(),
                                                ctypes.byref(chans[1]), ws.data_ptr(), ws.numel(), _stream()),
                   'scb_supcon_forward_peers')
        _lib.check(lib.scb_peer_wait(flags.data_ptr() + 64, state.data_ptr() + 8, 1, st.data_ptr() + 2 * n_k * 4,
                                     loss2.data_ptr(), plan.scalar_dtype, _stream()), 'scb_peer_wait')
        torch.cuda.synchronize()
        assert torch.equal(k, k_ref) and torch.equal(inv_norm, inv_ref)
        assert torch.allclose(st[0:n_k], stats[0], rtol=0, atol=0)
        assert torch.equal(st[n_k:2 * n_k], stats[2]) and torch.equal(possum, stats[1])
        assert float(loss2[0]) == float(loss_ref[0])
        assert float(L.typed_loss(loss2, dtype)) == float(L.typed_loss(loss_ref, dtype))
        assert state.tolist() == [0, it + 1, 0, it + 1] and int(flags[0]) == it + 1 and int(flags[16]) == it + 1


@pytest.mark.parametrize('labels', [False, True])
def test_both_row_layouts_give_the_reference_result(labels):
    """The same bf16 features as K in place (sample-major rows) and through prep (view-major rows)."""
    lib = _lib.load()
    dev = torch.device('cuda')
    B, V, D = 320, 2, 128
    rng = np.random.default_rng(11)
    z = F.normalize(torch.from_numpy(rng.standard_normal((B, V, D))), dim=-1).to(torch.bfloat16)
    lab = torch.from_numpy(rng.integers(0, 10, B)) if labels else None
    ref_loss, ref_grad, _ = supcon_forward_backward(z.double().numpy(), labels=None if lab is None else lab.numpy(),
                                                    temperature=0.5)
    feats = z.to(dev)
    lab_d = None if lab is None else lab.to(dev)
    results = []
    for layout in (_lib.ROWS_VIEW_MAJOR, _lib.ROWS_SAMPLE_MAJOR):
        plan = L.LossPlan(B, V, D, engine=_lib.ENGINE_BF16,
                          mask_kind=_lib.MASK_LABELS if labels else _lib.MASK_SIMCLR)
        plan.scalar_dtype = _lib.BF16
        plan.row_layout = layout
        if layout == _lib.ROWS_VIEW_MAJOR:
            k = torch.empty((plan.n_k, plan.ldk), dtype=torch.bfloat16, device=dev)
            L.launch_prep(lib, feats, plan, False, k, None, None)
        else:
            k = feats.view(plan.n_k, plan.ldk)
        prob = L._problem(plan, k, None, lab_d, None, 0.5, 0.07)
        stats = torch.empty((3, plan.n_k), dtype=torch.float32, device=dev)
        loss = torch.empty(2, dtype=torch.float32, device=dev)
        L.launch_forward(lib, prob, plan, stats[0], stats[1], stats[2], loss, dev)
        g = torch.ones((), dtype=torch.bfloat16, device=dev)
        dz = torch.empty_like(feats)
        L.launch_backward_features(lib, prob, plan, stats[0], stats[2], g, feats, False, None, dz, dev)
        torch.cuda.synchronize()
        results.append((float(loss[0]), dz.float().cpu().numpy()))
    tol = 2e-3 + 2.0 ** -8        # bf16 engine + bf16 rounding of the stored gradient
    for loss_v, grad in results:
        assert abs(loss_v - ref_loss) <= 2e-3 * abs(ref_loss)
        assert np.linalg.norm((grad - ref_grad).ravel()) <= tol * np.linalg.norm(ref_grad.ravel())
    assert abs(results[0][0] - results[1][0]) <= 1e-5 * abs(ref_loss)
