"""Bring-up check for the tensor-core loss engine: compares it with the fp32 engine piece by piece
(row statistics of the forward, then dK of the backward) on a few shapes and prints the errors.
Run on a GPU box:  python tools/tc_check.py"""
import ctypes
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from simclr_pytorch_distributed_b200 import _lib, losses as L   # noqa: E402


def run(engine, z, labels, mode, T=0.5, Tb=0.07, only_fwd=False):
    lib = _lib.load()
    dev = z.device
    B, V, D = z.shape
    mk = _lib.MASK_LABELS if labels is not None else _lib.MASK_SIMCLR
    plan = L.LossPlan(B, V, D, mode=mode, mask_kind=mk, engine=L._ENGINES[engine])
    k = torch.zeros((plan.n_k, plan.ldk), dtype=plan.k_torch, device=dev)
    stats = torch.zeros((4, plan.n_k), dtype=torch.float32, device=dev)
    loss = torch.zeros(2, dtype=torch.float32, device=dev)
    L.launch_prep(lib, z, plan, False, k, None, None)
    prob = L._problem(plan, k, None, labels, None, T, Tb)
    L.launch_forward(lib, prob, plan, stats[0], stats[1], stats[2], loss, dev)
    torch.cuda.synchronize()
    if only_fwd:
        return loss, stats, None
    g = torch.ones((), dtype=torch.float32, device=dev)
    dk = torch.zeros((plan.rows_local, plan.ldk), dtype=torch.float32, device=dev)
    L.launch_backward(lib, prob, plan, stats[0], stats[2], g, dk, dev)
    torch.cuda.synchronize()
    return loss, stats, dk[:, :D]


def main():
    torch.manual_seed(0)
    dev = torch.device('cuda')
    print('cc', _lib.load().scb_device_cc())
    cases = [(64, 2, 64, None, 0), (128, 2, 128, None, 0), (256, 2, 128, None, 0), (300, 2, 100, None, 0),
             (1024, 2, 128, None, 0), (200, 3, 128, 'lab', 0), (256, 2, 128, 'lab', 1), (256, 2, 128, None, 1),
             (4096, 2, 128, None, 0)]
    only = os.environ.get('TC_ONLY_FWD') == '1'
    for B, V, D, lab, mode in cases:
        z = F.normalize(torch.randn(B, V, D, device=dev) + 0.5 * torch.randn(1, 1, D, device=dev), dim=-1)
        z = z.to(torch.bfloat16)
        labels = torch.randint(0, max(2, B // 8), (B,), device=dev) if lab else None
        l0, s0, d0 = run('fp32', z.float(), labels, mode, only_fwd=only)
        l1, s1, d1 = run('bf16', z, labels, mode, only_fwd=only)
        n_a = B * V if mode == 0 else B
        e_lse = (s0[0, :n_a] - s1[0, :n_a]).abs().max().item()
        e_ps = (s0[1, :n_a] - s1[1, :n_a]).abs().max().item()
        e_np = (s0[2, :n_a] - s1[2, :n_a]).abs().max().item()
        msg = 'B=%d V=%d D=%d lab=%s mode=%d: loss %.6f vs %.6f | max|dlse| %.2e |dpossum| %.2e |dnpos| %.2e' % (
            B, V, D, bool(lab), mode, l0[0].item(), l1[0].item(), e_lse, e_ps, e_np)
        if d0 is not None:
            rel = ((d0 - d1).norm() / d0.norm()).item()
            msg += ' | dK rel-F %.2e max %.2e (ref max %.2e)' % (rel, (d0 - d1).abs().max().item(), d0.abs().max().item())
        print(msg, flush=True)


if __name__ == '__main__':
    main()
