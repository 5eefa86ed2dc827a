"""Row-sharded global-batch loss: the W-rank replacement of the reference's gather / normalise
glue plus criterion call, /root/reference/main_supcon.py:266-290.

The reference all-gathers the embeddings and then lets EVERY rank compute the full 2N x 2N loss
(and is hard-wired to 2 ranks, main_supcon.py:271-279).  Here each rank

  1. normalises its own rows (fused `F.normalize`, main_supcon.py:283) and all-gathers them (NCCL),
  2. runs the forward for ITS row block against all columns,
  3. all-gathers the per-row (log-sum-exp, positive count) and its share of the loss,

and returns the same global loss on every rank.  The backward needs no collective: with the
symmetric form of the gradient a rank computes exactly
d L_global / d (its own embeddings) -- what the reference obtains by re-inserting its local slice
into the gathered list (main_supcon.py:271-274) -- so DDP's parameter averaging works unchanged.

Row order of the gathered matrix: rank-major, view-major inside a rank (include/simclr_b200.h).
The loss is invariant under a consistent re-ordering of the samples, so this matches the reference's
[view 0 of every rank ; view 1 of every rank] ordering (main_supcon.py:276-279).
"""
from __future__ import annotations

import ctypes
import os
import warnings

import torch
import torch.distributed as dist
import torch.nn as nn

from . import _lib
from . import losses as L


class _CudaOps:
    """The four launches of the single-GPU path, re-used on the sharded problem."""
    requires_cuda = True
    prep = staticmethod(L.launch_prep)
    forward = staticmethod(L.launch_forward)
    backward = staticmethod(L.launch_backward)
    finish = staticmethod(L.launch_finish)
    backward_features = staticmethod(L.launch_backward_features)

    @staticmethod
    def lib():
        return _lib.load()


OPS = _CudaOps      # tests replace this with an oracle-backed stand-in to exercise the host logic on CPU


def _all_gather_flat(out, inp, group):
    """all_gather_into_tensor, or the list form on backends without it (gloo)."""
    if dist.get_backend(group) != 'gloo':
        dist.all_gather_into_tensor(out, inp, group=group)
    else:
        world = dist.get_world_size(group)
        parts = list(out.view(world, -1).unbind(0))
        dist.all_gather(parts, inp.reshape(-1), group=group)


class _PeerExchange:
    """The two exchanges of the sharded loss over peer memory instead of NCCL.

    Every rank owns symmetric-memory buffers (torch.distributed._symmetric_memory: plumbing only) that its
    peers map over NVLink.  Default: one kernel of this library (`scb_peer_bcast`) stores the rank's slice
    straight into every rank's buffer, followed by the handle's cross-rank stream barrier (8 us per exchange at
    2 GPUs against 12-18 us for the NCCL all-gather of the same few hundred kilobytes).  SCB_P2P=fused moves
    the stores into the PRODUCING kernels, which then raise per-rank flags (`scb_supcon_prep_peers`,
    `scb_supcon_forward_peers`, consumer side `scb_peer_wait`); it is correct but measured slower so far
    (15 us per exchange: flag latency at system scope), so it is not the default.  SCB_P2P=0: NCCL all-gathers.

    Buffers rotate through GENERATIONS copies: a rank may only overwrite its peers' copy g once they are
    done reading it, and the barrier of the generations in between guarantees that (each forward is
    followed in stream order by its backward).  A backward that finds its generation re-used raises.
    """
    GENERATIONS = 4
    _cache = {}
    _failed = False

    def __init__(self, group, dev, n_k, ldk, k_dtype, rows, world, rank):
        import torch.distributed._symmetric_memory as symm
        self.world, self.rank, self.rows, self.n_k, self.ldk = world, rank, rows, n_k, ldk
        self.k_bufs, self.st_bufs, self.k_ptrs, self.st_ptrs, self.handles = [], [], [], [], []
        self.st_len = 2 * n_k + 4 * world                      # [lse(n_k) | npos(n_k) | loss share per rank, padded to 4]
        for _ in range(self.GENERATIONS):
            kb = symm.empty((n_k, ldk), dtype=k_dtype, device=dev)
            kh = symm.rendezvous(kb, group)
            sb = symm.empty((self.st_len,), dtype=torch.float32, device=dev)
            sh = symm.rendezvous(sb, group)
            self.k_bufs.append(kb)
            self.st_bufs.append(sb)
            self.k_ptrs.append((ctypes.c_void_p * world)(*[int(p) for p in kh.buffer_ptrs]))
            self.st_ptrs.append((ctypes.c_void_p * world)(*[int(p) for p in sh.buffer_ptrs]))
            self.handles.append((kh, sh))
        # fused mode: flag words [2 channels][16 ranks] in symmetric memory, local ticket/counter per channel
        fl = symm.empty((32,), dtype=torch.int32, device=dev)
        fl.zero_()
        fh = symm.rendezvous(fl, group)
        torch.cuda.synchronize(dev)
        fh.barrier(channel=0)                 # nobody raises a flag before everybody has cleared its own
        torch.cuda.synchronize(dev)
        self.flags, self.flags_handle = fl, fh
        self.state = torch.zeros(4, dtype=torch.int32, device=dev)
        self.channels = []                    # [generation][channel] -> scb_peer_channel
        for g in range(self.GENERATIONS):
            pair = []
            for c, ptrs in ((0, self.k_ptrs[g]), (1, self.st_ptrs[g])):
                ch = _lib.PeerChannel()
                for w in range(world):
                    ch.dst[w] = ptrs[w]
                    ch.flag[w] = int(fh.buffer_ptrs[w]) + (c * 16 + rank) * 4
                ch.state = self.state.data_ptr() + c * 8
                ch.n_dst = world
                pair.append(ch)
            self.channels.append(pair)
        self.calls = 0
        self.capture_id, self.capture_calls = 0, 0
        esz = kb.element_size()
        self.k_seg = (ctypes.c_int64 * 3)(0, rank * rows * ldk * esz, rows * ldk * esz)
        self.st_seg = (ctypes.c_int64 * 9)(0, rank * rows * 4, rows * 4,
                                           rows * 4, (n_k + rank * rows) * 4, rows * 4,
                                           2 * rows * 4, (2 * n_k + 4 * rank) * 4, 16)

    @classmethod
    def get(cls, group, dev, n_k, ldk, k_dtype, rows, world, rank):
        """The exchange for this problem shape, or None when peer memory is not to be used."""
        if cls._failed or os.environ.get('SCB_P2P', '1') == '0' or dist.get_backend(group) != 'nccl' or world > 16:
            return None
        if (rows * ldk * torch.empty((), dtype=k_dtype).element_size()) % 16 or (rows * 4) % 16:
            return None
        key = (id(group), dev.index, n_k, ldk, k_dtype, rows, world)
        if key not in cls._cache:
            try:
                cls._cache[key] = cls(group, dev, n_k, ldk, k_dtype, rows, world, rank)
            except Exception as e:       # no symmetric memory on this system: every rank fails alike -> NCCL
                cls._failed = True
                warnings.warn('peer-memory exchange unavailable (%s); using NCCL all-gather' % (e,))
                return None
        return cls._cache[key]

    def next_generation(self, lib, dev):
        """Buffer generation of the next forward, and whether the first exchange must be preceded by a barrier.

        Eager calls rotate through the generations.  Inside a CUDA-graph capture the choice is baked into the graph
        and replays repeat it, so the rotation restarts at every capture (all ranks capture alike) and the first
        exchange of the graph waits for every rank to have finished whatever ran before: the seam between two
        replays -- or between eager steps and a replay -- can then never hand a rank a buffer that a peer is
        still reading, whatever the number of steps in the graph."""
        cap = lib.scb_stream_capture_id(L._stream(dev)) if dev.type == 'cuda' else 0
        pre_barrier = False
        if cap > 0:
            if cap != self.capture_id:
                self.capture_id, self.capture_calls, pre_barrier = cap, 0, True
            g = self.capture_calls % self.GENERATIONS
            self.capture_calls += 1
            self.calls += 1
            return g, self.calls, pre_barrier
        if self.capture_id != 0:           # first eager call after replays of a captured graph
            self.capture_id, pre_barrier = 0, True
        g = self.calls % self.GENERATIONS
        self.calls += 1
        return g, self.calls, pre_barrier

    def check_alive(self, call_index):
        if self.calls - call_index >= self.GENERATIONS:
            raise RuntimeError('DistributedSupConLoss: backward of a forward that is more than %d calls old '
                               '(its gathered embeddings were recycled)' % (self.GENERATIONS - 1))

    def _bcast(self, lib, src, seg, n_seg, ptrs, handle, dev):
        _lib.check(lib.scb_peer_bcast(src.data_ptr(), seg, n_seg, ptrs, self.world, L._stream(dev)), 'scb_peer_bcast')
        handle.barrier(channel=0)

    def wait(self, lib, channel, dev, loss_shares=None, out_loss=None, scalar_dtype=0):
        """Consumer side of a fused exchange: the stream blocks until every rank's flag shows it."""
        _lib.check(lib.scb_peer_wait(self.flags.data_ptr() + channel * 64, self.state.data_ptr() + channel * 8,
                                     self.world, loss_shares, out_loss, scalar_dtype, L._stream(dev)), 'scb_peer_wait')

    def share_rows(self, lib, g, k_local, dev):
        self._bcast(lib, k_local, self.k_seg, 1, self.k_ptrs[g], self.handles[g][0], dev)
        return self.k_bufs[g]

    def share_stats(self, lib, g, packed, dev):
        self._bcast(lib, packed, self.st_seg, 3, self.st_ptrs[g], self.handles[g][1], dev)
        return self.st_bufs[g]


class _DistSupConFunction(torch.autograd.Function):

    @staticmethod
    def forward(ctx, feats, labels_all, plan, temperature, base_temperature, normalize, group):
        ops = OPS
        lib = ops.lib()
        dev = feats.device
        rows, n_k, ldk = plan.rows_local, plan.n_k, plan.ldk
        world = plan.n_seg
        inv_norm = torch.empty(rows, dtype=torch.float32, device=dev) if normalize else None
        px = (_PeerExchange.get(group, dev, n_k, ldk, plan.k_torch, rows, world, plan.seg_index)
              if ops.requires_cuda else None)
        gen, call_index, pre_barrier = px.next_generation(lib, dev) if px is not None else (0, 0, False)
        if pre_barrier:
            px.handles[gen][0].barrier(channel=0)
        fused = px is not None and os.environ.get('SCB_P2P', '1') == 'fused'
        possum = torch.empty(rows, dtype=torch.float32, device=dev)
        if fused:
            # exchange 1: the prep kernel stores this rank's rows into every rank's gathered matrix
            _lib.check(lib.scb_supcon_prep_peers(
                feats.data_ptr(), _lib.dtype_code(feats.dtype), plan.n_local, plan.n_views, plan.dim,
                feats.stride(0), feats.stride(1), int(normalize), 1e-12, ctypes.byref(px.channels[gen][0]),
                plan.row_begin, plan.k_dtype, ldk, inv_norm.data_ptr() if inv_norm is not None else None,
                L._stream(dev)), 'scb_supcon_prep_peers')
            px.wait(lib, 0, dev)
            k_all = px.k_bufs[gen]
            prob = L._problem(plan, k_all, None, labels_all, None, temperature, base_temperature)
            # exchange 2: the finalize kernel stores (lse, npos, loss share) into every rank's statistics array
            need = lib.scb_supcon_workspace_bytes(ctypes.byref(prob), rows)
            ws = L._Workspace.get(dev, need)
            _lib.check(lib.scb_supcon_forward_peers(
                ctypes.byref(prob), plan.row_begin, rows, plan.n_anchor_total, possum.data_ptr(),
                ctypes.byref(px.channels[gen][1]), ws.data_ptr(), ws.numel(), L._stream(dev)), 'scb_supcon_forward_peers')
            st = px.st_bufs[gen]
            loss2 = torch.empty(2, dtype=torch.float32, device=dev)
            px.wait(lib, 1, dev, st.data_ptr() + 2 * n_k * 4, loss2.data_ptr(), plan.scalar_dtype)
            lse_all, npos_all = st[0:n_k], st[n_k:2 * n_k]
            ctx.px, ctx.call_index = px, call_index
            ctx.plan, ctx.args = plan, (temperature, base_temperature, normalize)
            ctx.save_for_backward(feats, labels_all, k_all, lse_all, npos_all, inv_norm)
            return L.typed_loss(loss2, feats.dtype)
        # bf16 features that need no normalisation are this rank's rows already (sample-major order inside the
        # segment, include/simclr_b200.h): no stacking copy, the exchange reads them where they are
        # The choice (and with it the row order of every segment of the gathered matrix) depends only on things that
        # are the same on every rank; a rank whose tensor is strided or misaligned copies it, keeping the order.
        zero_copy = (ops.requires_cuda and plan.engine == _lib.ENGINE_BF16 and feats.dtype == torch.bfloat16 and
                     not normalize and plan.mode == _lib.MODE_ALL and plan.dim == ldk)
        if zero_copy:
            plan.row_layout = _lib.ROWS_SAMPLE_MAJOR
            src = feats.detach()
            if not src.is_contiguous() or src.data_ptr() % 16 != 0:
                src = src.contiguous().clone()
            k_local = src.view(rows, ldk)
        else:
            k_local = torch.empty((rows, ldk), dtype=plan.k_torch, device=dev)
            ops.prep(lib, feats, plan, normalize, k_local, None, inv_norm)
        if px is not None:
            k_all = px.share_rows(lib, gen, k_local, dev)                        # exchange 1 over peer memory
        else:
            k_all = torch.empty((n_k, ldk), dtype=plan.k_torch, device=dev)
            _all_gather_flat(k_all, k_local, group)                              # exchange 1: embeddings
        # local row statistics: [lse(rows) | npos(rows) | loss share, 3 pad]
        packed = torch.empty(2 * rows + 4, dtype=torch.float32, device=dev)
        prob = L._problem(plan, k_all, None, labels_all, None, temperature, base_temperature)
        ops.forward(lib, prob, plan, packed[0:rows], possum, packed[rows:2 * rows], packed[2 * rows:2 * rows + 1], dev)
        if px is not None:
            st = px.share_stats(lib, gen, packed, dev)                           # exchange 2 over peer memory
            lse_all, npos_all = st[0:n_k], st[n_k:2 * n_k]
            loss = st[2 * n_k:2 * n_k + 4 * world:4].sum()
        else:
            gathered = torch.empty((world, 2 * rows + 4), dtype=torch.float32, device=dev)
            _all_gather_flat(gathered, packed, group)                           # exchange 2: row statistics
            lse_all = gathered[:, 0:rows].contiguous().view(-1)
            npos_all = gathered[:, rows:2 * rows].contiguous().view(-1)
            loss = gathered[:, 2 * rows].sum()       # every rank's share of mean_i loss_i (losses.py:91)
        ctx.px, ctx.call_index = px, call_index
        ctx.plan, ctx.args = plan, (temperature, base_temperature, normalize)
        ctx.save_for_backward(feats, labels_all, k_all, lse_all, npos_all, inv_norm)
        return loss.to(feats.dtype) if feats.dtype != torch.float32 else loss

    @staticmethod
    def backward(ctx, grad_out):
        ops = OPS
        lib = ops.lib()
        feats, labels_all, k_all, lse_all, npos_all, inv_norm = ctx.saved_tensors
        if ctx.px is not None:
            ctx.px.check_alive(ctx.call_index)
        plan = ctx.plan
        temperature, base_temperature, normalize = ctx.args
        dev = feats.device
        g = L.grad_scalar(grad_out, feats.dtype)
        prob = L._problem(plan, k_all, None, labels_all, None, temperature, base_temperature)
        dfeats = torch.empty(feats.shape, dtype=feats.dtype, device=dev)
        if plan.engine != _lib.ENGINE_FP32 and hasattr(ops, 'backward_features'):
            ops.backward_features(lib, prob, plan, lse_all, npos_all, g, feats, normalize, inv_norm, dfeats, dev)
        else:                                                                    # no collective needed either way
            dk = torch.empty((plan.rows_local, plan.ldk), dtype=torch.float32, device=dev)
            ops.backward(lib, prob, plan, lse_all, npos_all, g, dk, dev)
            ops.finish(lib, dk, feats, plan, normalize, inv_norm, dfeats)
        return dfeats, None, None, None, None, None, None


class DistributedSupConLoss(nn.Module):
    """SupCon / NT-Xent over the global batch with the rows sharded across the ranks.

    forward(features, labels=None): `features` [bsz, n_views, ...] are THIS rank's embeddings (raw
    model outputs; `normalize=True` applies main_supcon.py:283), `labels` [bsz] this rank's labels.
    Every rank must pass the same bsz and n_views.  Returns the global loss (identical on all ranks);
    its gradient with respect to `features` is d L_global / d features, as in the reference.
    Explicit `mask` arguments are a single-GPU feature of `SupConLoss` (the reference cannot shard
    them either).
    """

    def __init__(self, temperature=0.07, contrast_mode='all', base_temperature=0.07, engine='auto',
                 normalize=True, process_group=None, graph=False):
        super().__init__()
        # graph=True: as SupConLoss(graph=True) -- forward + backward (both exchanges included) captured once per shape
        # and replayed; the input is always copied into a buffer the module owns, so that "capture or replay" depends
        # on the shape alone and every rank decides alike.  SimCLR mode only (labels take the eager path).
        self.graph = graph
        self._graph_cache, self._graph_cache_settings, self._graph_hot = None, None, None
        self.temperature = temperature
        self.contrast_mode = contrast_mode
        self.base_temperature = base_temperature
        self.engine = engine
        self.normalize = normalize
        self.process_group = process_group

    def _graph_settings(self):
        return (self.temperature, self.contrast_mode, self.base_temperature, self.engine, self.normalize,
                self.process_group)

    def forward(self, features, labels=None):
        hot = self._graph_hot
        if (hot is not None and labels is None and self.graph and features.shape == hot[0] and
                features.stride() == hot[1] and features.dtype is hot[2] and self._graph_settings() == hot[3] and
                features.requires_grad and torch.is_grad_enabled() and not torch.cuda.is_current_stream_capturing()):
            hot[4].launch(features)        # same shape as the last graphed call: replay; the GPU starts before
            return L._GraphedSupCon.apply(features, hot[4], True)      # autograd's bookkeeping for the call
        labels, mode = L.validate_inputs(features, labels, None, self.contrast_mode)
        if OPS.requires_cuda and not features.is_cuda:
            raise RuntimeError('DistributedSupConLoss runs on CUDA tensors only (there is no CPU fallback); '
                               'got a %s tensor' % features.device.type)
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError('DistributedSupConLoss needs an initialised torch.distributed process group; '
                               'use SupConLoss on a single GPU')
        group = self.process_group if self.process_group is not None else dist.group.WORLD
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        if len(features.shape) > 3:
            features = features.view(features.shape[0], features.shape[1], -1)
        if features.stride(2) != 1 and features.shape[2] > 1:
            features = features.contiguous()
        bsz, n_views, dim = features.shape
        dev = features.device
        if (self.graph and labels is None and OPS.requires_cuda and features.requires_grad and torch.is_grad_enabled()
                and not torch.cuda.is_current_stream_capturing()):
            settings = self._graph_settings()
            if self._graph_cache is None or self._graph_cache_settings != settings:   # baked into the captured graphs
                self._graph_cache = L.GraphCache(
                    lambda: DistributedSupConLoss(self.temperature, self.contrast_mode, self.base_temperature,
                                                  self.engine, self.normalize, self.process_group),
                    bind_addresses=False)
                self._graph_cache_settings = settings
            out = self._graph_cache.run(features, None)
            self._graph_hot = (features.shape, features.stride(), features.dtype, settings, self._graph_cache.last_entry)
            return out
        labels_all = None
        mask_kind = _lib.MASK_SIMCLR
        if labels is not None:
            mask_kind = _lib.MASK_LABELS
            labels = L.label_ids(labels, dev)
            labels_all = torch.empty(world * bsz, dtype=torch.int64, device=dev)
            _all_gather_flat(labels_all, labels, group)       # sample id = rank*bsz + b
        engine = L.select_engine(self.engine, features.dtype, dim, mask_kind, world * n_views * bsz, allow_x3=False)
        if engine == _lib.ENGINE_BF16X3:
            raise ValueError("engine 'bf16x3' is a single-GPU engine of SupConLoss; the row-sharded loss takes 'fp32' or 'bf16'")
        plan = L.LossPlan(bsz, n_views, dim, n_seg=world, seg_index=rank, mode=mode, mask_kind=mask_kind,
                          engine=engine)
        plan.scalar_dtype = _lib.dtype_code(features.dtype)
        return _DistSupConFunction.apply(features, labels_all, plan, float(self.temperature),
                                         float(self.base_temperature), bool(self.normalize), group)
