"""Where the time of one row-sharded loss step goes (diagnostics).  Each section is captured 20x in a CUDA
graph and replayed, so launch latencies are the in-graph ones.
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/prof_dist.py"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from simclr_pytorch_distributed_b200 import _lib, losses as L, gather as G   # noqa: E402

rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
torch.cuda.set_device(int(os.environ['LOCAL_RANK']))
dev = torch.device('cuda', int(os.environ['LOCAL_RANK']))
dist.init_process_group('nccl', device_id=dev)
lib = _lib.load()
group = dist.group.WORLD
B_glob, V, D = 4096, 2, 128
bsz = B_glob // world
feats = torch.nn.functional.normalize(torch.randn(bsz, V, D, device=dev), dim=-1).to(torch.bfloat16)
plan = L.LossPlan(bsz, V, D, n_seg=world, seg_index=rank, engine=_lib.ENGINE_BF16)
plan.scalar_dtype = _lib.BF16
rows, n_k, ldk = plan.rows_local, plan.n_k, plan.ldk
k_local = torch.empty((rows, ldk), dtype=torch.bfloat16, device=dev)
px = G._PeerExchange.get(group, dev, n_k, ldk, torch.bfloat16, rows, world, rank)
assert px is not None
packed = torch.empty(2 * rows + 4, dtype=torch.float32, device=dev)
possum = torch.empty(rows, dtype=torch.float32, device=dev)
gout = torch.ones((), dtype=torch.bfloat16, device=dev)
dfeats = torch.empty_like(feats)
k_all = px.k_bufs[0]
st = px.st_bufs[0]
prob = L._problem(plan, k_all, None, None, None, 0.5, 0.07)
kn = torch.empty((n_k, ldk), dtype=torch.bfloat16, device=dev)
gathered = torch.empty((world, 2 * rows + 4), dtype=torch.float32, device=dev)


def sec_prep():
    L.launch_prep(lib, feats, plan, False, k_local, None, None)


def sec_share_rows():
    px.share_rows(lib, 0, k_local, dev)


def sec_bcast_only():
    _lib.check(lib.scb_peer_bcast(k_local.data_ptr(), px.k_seg, 1, px.k_ptrs[0], world, L._stream(dev)), 'bcast')


def sec_barrier_only():
    px.handles[0][0].barrier(channel=0)


def sec_forward():
    L.launch_forward(lib, prob, plan, packed[0:rows], possum, packed[rows:2 * rows], packed[2 * rows:2 * rows + 1], dev)


def sec_share_stats():
    px.share_stats(lib, 0, packed, dev)


def sec_glue():
    return st[2 * n_k:2 * n_k + 4 * world:4].sum().to(torch.bfloat16)


def sec_backward():
    L.launch_backward_features(lib, prob, plan, st[0:n_k], st[n_k:2 * n_k], gout, feats, False, None, dfeats, dev)


def sec_nccl_rows():
    dist.all_gather_into_tensor(kn, k_local, group=group)


def sec_nccl_stats():
    dist.all_gather_into_tensor(gathered, packed, group=group)


import ctypes   # noqa: E402


def sec_prep_peers():
    _lib.check(lib.scb_supcon_prep_peers(feats.data_ptr(), _lib.BF16, bsz, V, D, feats.stride(0), feats.stride(1), 0, 1e-12,
                                         ctypes.byref(px.channels[0][0]), plan.row_begin, plan.k_dtype, ldk, None,
                                         L._stream(dev)), 'prep_peers')
    px.wait(lib, 0, dev)


ws_need = lib.scb_supcon_workspace_bytes(ctypes.byref(prob), rows)
loss2 = torch.empty(2, dtype=torch.float32, device=dev)


def sec_forward_peers():
    ws = L._Workspace.get(dev, ws_need)
    _lib.check(lib.scb_supcon_forward_peers(ctypes.byref(prob), plan.row_begin, rows, plan.n_anchor_total, possum.data_ptr(),
                                            ctypes.byref(px.channels[0][1]), ws.data_ptr(), ws.numel(), L._stream(dev)),
               'forward_peers')
    px.wait(lib, 1, dev, st.data_ptr() + 2 * n_k * 4, loss2.data_ptr(), plan.scalar_dtype)


secs = [('fused: prep_peers + wait', sec_prep_peers), ('fused: forward_peers + wait(+loss)', sec_forward_peers),
        ('prep', sec_prep), ('share rows (bcast + barrier)', sec_share_rows), ('  bcast kernel alone', sec_bcast_only),
        ('  barrier alone', sec_barrier_only), ('forward + finalize', sec_forward),
        ('share stats (bcast + barrier)', sec_share_stats), ('loss glue (sum, cast)', sec_glue),
        ('backward + reduce/finish', sec_backward), ('NCCL all_gather rows', sec_nccl_rows),
        ('NCCL all_gather stats', sec_nccl_stats)]
for _n, f in secs:          # warm up everything once (rendezvous, lazy init)
    f()
    torch.cuda.synchronize()
    if rank == 0:
        print('warm ' + _n, flush=True)
torch.cuda.synchronize()
dist.barrier()
REP = 20
s = torch.cuda.Stream()
for name, f in secs:
    if rank == 0:
        print('.. ' + name, flush=True)
    with torch.cuda.stream(s):
        for _ in range(3):
            f()
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(REP):
                f()
        g.replay()
        s.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(5):
            g.replay()
        e1.record(s)
        s.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / (5 * REP)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print('%-34s %7.2f us' % (name, t.item()))
sys.stdout.flush()
os._exit(0)      # tearing the process group down with captured graphs alive can hang
