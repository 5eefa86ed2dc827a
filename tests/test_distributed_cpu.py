"""World-size > 1 host logic on CPU with the `gloo` backend (no GPU needed).

The product kernels only run on a GPU, so here the four loss launches and the four batch-norm
launches are replaced by oracle-backed stand-ins that honour the same C-ABI contract (they decode
the very `scb_supcon_problem` struct the CUDA library would get).  What is under test is everything
around the kernels: the all-gathers, the row order of the gathered matrix, which rows a rank owns,
the statistics exchange, the gradient slice each rank ends up with -- against golden vectors
produced by the unmodified reference glue (tests/golden/glue_w*.npz, bn_w*.npz)."""
import ctypes
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import golden


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _as_array(ptr, n, ctype, dtype):
    return np.ctypeslib.as_array((ctype * n).from_address(ptr)).view(dtype)


# ----------------------------------------------------------------------------- loss stand-ins
class OracleLossOps:
    """CPU stand-ins for scb_supcon_prep / forward / backward / finish (include/simclr_b200.h)."""
    requires_cuda = False

    @staticmethod
    def lib():
        return None

    @staticmethod
    def prep(lib, feats, plan, normalize, k_rows, k_lo, inv_norm):
        from oracle import normalize_rows
        f = feats.detach().double().numpy()
        rows = np.transpose(f, (1, 0, 2)).reshape(plan.rows_local, -1)        # row = v*n_local + b
        if normalize:
            norm = np.maximum(np.sqrt((rows * rows).sum(-1)), 1e-12)
            inv_norm.copy_(torch.from_numpy(1.0 / norm).float())
            rows = normalize_rows(rows)
        k_rows.zero_()
        k_rows[:, :plan.dim] = torch.from_numpy(rows).to(k_rows.dtype)

    @staticmethod
    def _decode(prob, plan):
        n_k, ldk, D = prob.n_k, prob.ldk, prob.dim
        K = _as_array(prob.k, n_k * ldk, ctypes.c_float, np.float32).reshape(n_k, ldk)[:, :D].astype(np.float64)
        V, seg, W = prob.n_views, prob.seg, prob.n_seg
        # internal row q*V*seg + v*seg + b  ->  reference features [sample q*seg+b, view v]
        feats_ref = K.reshape(W, V, seg, D).transpose(0, 2, 1, 3).reshape(W * seg, V, D)
        labels = None
        if prob.labels:
            labels = _as_array(prob.labels, W * seg, ctypes.c_int64, np.int64).copy()
        mode = 'all' if prob.mode == 0 else 'one'
        T = 1.0 / prob.inv_temperature
        # map: local row index (v*seg + b of segment q) -> reference row v*Bg + q*seg + b
        q = plan.seg_index
        v_idx, b_idx = np.divmod(np.arange(plan.rows_local), seg)
        ref_rows = v_idx * (W * seg) + q * seg + b_idx
        return feats_ref, labels, mode, T, ref_rows

    @staticmethod
    def forward(lib, prob, plan, lse, possum, npos, loss, dev):
        from oracle import supcon_row_stats
        feats_ref, labels, mode, T, ref_rows = OracleLossOps._decode(prob, plan)
        o_lse, o_ps, o_np = supcon_row_stats(feats_ref, labels, None, T, mode)
        A = len(o_lse)
        is_anchor = ref_rows < A
        rr = np.where(is_anchor, ref_rows, 0)
        lse.copy_(torch.from_numpy(np.where(is_anchor, o_lse[rr], 0.0)).float())
        possum.copy_(torch.from_numpy(np.where(is_anchor, o_ps[rr], 0.0)).float())
        npos.copy_(torch.from_numpy(np.where(is_anchor, o_np[rr], 1.0)).float())
        rows_loss = -prob.loss_scale * (o_ps[rr] / o_np[rr] - o_lse[rr])
        loss.copy_(torch.tensor([np.where(is_anchor, rows_loss, 0.0).sum() / plan.n_anchor_total]).float())

    @staticmethod
    def backward(lib, prob, plan, lse_all, npos_all, grad_out, dk, dev):
        from oracle import supcon_forward_backward
        feats_ref, labels, mode, T, ref_rows = OracleLossOps._decode(prob, plan)
        Tb = T / prob.loss_scale
        _, dfeat, _ = supcon_forward_backward(feats_ref, labels, None, T, mode, Tb, grad_out=float(grad_out))
        Bg, V, D = feats_ref.shape
        dK_ref = np.transpose(dfeat, (1, 0, 2)).reshape(V * Bg, D)
        dk.zero_()
        dk[:, :D] = torch.from_numpy(dK_ref[ref_rows]).float()

    @staticmethod
    def finish(lib, dk, feats, plan, normalize, inv_norm, dfeats):
        from oracle import normalize_rows_backward
        g = dk[:, :plan.dim].double().numpy()
        f = feats.detach().double().numpy()
        rows = np.transpose(f, (1, 0, 2)).reshape(plan.rows_local, -1)
        if normalize:
            g = normalize_rows_backward(rows, g)
        out = g.reshape(plan.n_views, plan.n_local, -1).transpose(1, 0, 2)
        dfeats.copy_(torch.from_numpy(np.ascontiguousarray(out)).to(dfeats.dtype))


def _loss_worker(rank, world, port, fixture, q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from simclr_pytorch_distributed_b200 import gather
        gather.OPS = OracleLossOps
        g = golden(fixture)
        bsz = int(g['bsz'])
        part = torch.from_numpy(g[f'part{rank}']).float()                     # [2*bsz, D], view-major
        feats = part.view(2, bsz, -1).transpose(0, 1).contiguous().requires_grad_(True)   # [bsz, 2, D]
        crit = gather.DistributedSupConLoss(temperature=float(g['temperature']), engine='fp32')
        loss = crit(feats)
        loss.backward()
        grad = feats.grad.transpose(0, 1).reshape(2 * bsz, -1).numpy()       # back to the part layout
        ok_loss = abs(float(loss) - float(g['loss'])) <= 1e-5 * abs(float(g['loss']))
        ref = g[f'grad{rank}']
        ok_grad = np.abs(grad - ref).max() <= 1e-5 * np.abs(ref).max()
        q.put((rank, bool(ok_loss), bool(ok_grad), float(loss)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('fixture,world', [('glue_w2', 2), ('glue_w4', 4)])
def test_row_sharded_loss_matches_reference_glue(fixture, world):
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_loss_worker, args=(r, world, port, fixture, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    results = sorted(q.get(timeout=5) for _ in range(world))
    assert [r[0] for r in results] == list(range(world))
    assert all(r[1] for r in results), results          # same global loss on every rank
    assert all(r[2] for r in results), results          # each rank gets d L_global / d its own slice


def _labels_worker(rank, world, port, q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from simclr_pytorch_distributed_b200 import gather
        from oracle import supcon_forward_backward
        gather.OPS = OracleLossOps
        rng = np.random.default_rng(77)
        bsz, V, D = 6, 2, 8
        z_all = rng.standard_normal((world * bsz, V, D))
        z_all /= np.linalg.norm(z_all, axis=-1, keepdims=True)
        lab_all = rng.integers(0, 3, size=world * bsz)
        sl = slice(rank * bsz, (rank + 1) * bsz)
        for mode in ('all', 'one'):
            f = torch.from_numpy(z_all[sl]).float().requires_grad_(True)
            crit = gather.DistributedSupConLoss(temperature=0.2, contrast_mode=mode, engine='fp32', normalize=False)
            loss = crit(f, torch.from_numpy(lab_all[sl]))
            loss.backward()
            ref_loss, ref_grad, _ = supcon_forward_backward(z_all, lab_all, None, 0.2, mode)
            ok = abs(float(loss) - ref_loss) <= 1e-5 * abs(ref_loss) and \
                np.abs(f.grad.numpy() - ref_grad[sl]).max() <= 1e-5 * np.abs(ref_grad).max()
            q.put((rank, mode, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_row_sharded_loss_with_labels_and_mode_one():
    world = 2
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_labels_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    results = [q.get(timeout=5) for _ in range(2 * world)]
    assert all(r[2] for r in results), results


# ----------------------------------------------------------------------------- SyncBN stand-ins
class OracleBNKernels:
    """CPU stand-ins for scb_bn_stats / apply / backward_reduce / backward_apply."""

    @staticmethod
    def stats(x, geom, out_stats):
        xd = x.detach().double()
        C = geom[2]
        ax = [a for a in range(xd.dim()) if a != 1]
        mean = xd.mean(dim=ax)
        m2 = ((xd - mean.view(1, C, *([1] * (xd.dim() - 2)))) ** 2).sum(dim=ax)
        out_stats[:C] = mean.float()
        out_stats[C:2 * C] = m2.float()
        out_stats[2 * C] = float(xd.numel() // C)

    @staticmethod
    def apply(x, y, geom, stats_all, n_ranks, weight, bias, running_mean, running_var, momentum, eps, saved, flags,
              nbt=None):
        C = geom[2]
        if nbt is not None:
            nbt.add_(1)                                # the kernel counts the step (batchnorm.py:763-764)
        st = stats_all.double().view(n_ranks, 2 * C + 1)
        n, mean, m2 = 0.0, torch.zeros(C, dtype=torch.float64), torch.zeros(C, dtype=torch.float64)
        for r in range(n_ranks):                       # Chan merge with counts, rank order
            nr = float(st[r, 2 * C])
            if nr < 1:
                continue
            delta = st[r, :C] - mean
            tot = n + nr
            mean = mean + delta * (nr / tot)
            m2 = m2 + st[r, C:2 * C] + delta * delta * (n * nr / tot)
            n = tot
        invstd = 1.0 / torch.sqrt(m2 / n + eps)
        running_mean.mul_(1 - momentum).add_(momentum * mean.float())
        running_var.mul_(1 - momentum).add_(momentum * (m2 / (n - 1)).float())
        saved[:C] = mean.float()
        saved[C:2 * C] = invstd.float()
        saved[2 * C] = n
        shape = (1, C) + (1,) * (x.dim() - 2)
        out = (x.double() - mean.view(shape)) * invstd.view(shape) * weight.double().view(shape) + bias.double().view(shape)
        y.copy_(out.to(y.dtype))

    @staticmethod
    def backward_reduce(dy, x, geom, saved, weight, bias, sums, grad_weight, grad_bias, flags):
        C = geom[2]
        shape = (1, C) + (1,) * (x.dim() - 2)
        ax = [a for a in range(x.dim()) if a != 1]
        mean, invstd = saved[:C].double(), saved[C:2 * C].double()
        s1 = dy.double().sum(dim=ax)
        s2 = (dy.double() * (x.double() - mean.view(shape))).sum(dim=ax)
        sums[:C] = s1.float()
        sums[C:] = s2.float()
        grad_weight.copy_((s2 * invstd).float())
        grad_bias.copy_(s1.float())

    @staticmethod
    def backward_apply(dy, x, dx, geom, saved, weight, bias, sums, flags):
        C = geom[2]
        shape = (1, C) + (1,) * (x.dim() - 2)
        mean, invstd, n = saved[:C].double(), saved[C:2 * C].double(), float(saved[2 * C])
        s1, s2 = sums[:C].double(), sums[C:].double()
        out = (dy.double() - (s1 / n).view(shape) - (x.double() - mean.view(shape)) * (invstd * invstd * s2 / n).view(shape)) \
            * (weight.double() * invstd).view(shape)
        dx.copy_(out.to(dx.dtype))


def _bn_worker(rank, world, port, fixture, q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from simclr_pytorch_distributed_b200.sync_batchnorm import _SyncBNFunction
        _SyncBNFunction.kernels = OracleBNKernels
        g = golden(fixture)
        x = torch.from_numpy(g[f'x{rank}']).float().requires_grad_(True)
        dy = torch.from_numpy(g[f'dy{rank}']).float()
        w = torch.from_numpy(g['weight']).float().requires_grad_(True)
        b = torch.from_numpy(g['bias']).float().requires_grad_(True)
        rm = torch.from_numpy(g['running_mean0']).float()
        rv = torch.from_numpy(g['running_var0']).float()
        y = _SyncBNFunction.apply(x, None, w, b, rm, rv, float(g['eps']), float(g['momentum']), dist.group.WORLD,
                                  world, False, True)
        y.backward(dy)
        tol = 2e-5
        ok = (np.abs(y.detach().numpy() - g[f'y{rank}']).max() <= tol * max(1, np.abs(g[f'y{rank}']).max())
              and np.abs(x.grad.numpy() - g[f'dx{rank}']).max() <= tol * max(1, np.abs(g[f'dx{rank}']).max())
              and np.abs(rm.numpy() - g['running_mean1']).max() <= tol
              and np.abs(rv.numpy() - g['running_var1']).max() <= tol)
        # dgamma / dbeta stay per-rank (DDP averages them later): their sum is the global gradient
        gw = w.grad.clone()
        dist.all_reduce(gw)
        ok = ok and np.abs(gw.numpy() - g['grad_weight']).max() <= 1e-4 * max(1, np.abs(g['grad_weight']).max())
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('fixture,world', [('bn_w2', 2), ('bn_w3_ragged', 3)])
def test_syncbn_exchange_matches_concatenated_batch(fixture, world):
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_bn_worker, args=(r, world, port, fixture, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    results = sorted(q.get(timeout=5) for _ in range(world))
    assert all(r[1] for r in results), results
