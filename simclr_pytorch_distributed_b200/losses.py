"""`SupConLoss` -- drop-in for /root/reference/losses.py:7-93 on a B200.

Same constructor, same `forward(features, labels=None, mask=None)`, same `ValueError`s, same
value and gradient; the arithmetic runs in hand-written sm_100a kernels behind the C ABI of
include/simclr_b200.h and never materialises the [A, V*B] similarity matrix.  CUDA tensors
only: there is no CPU path (the reference's own code is the CPU path).

Extras (all default to the reference's behaviour):
  engine     'auto' | 'fp32' | 'bf16' | 'bf16x3'   arithmetic of the similarity tiles
  normalize  fuse `F.normalize(features, dim=-1)` (main_supcon.py:283) into the kernels
"""
from __future__ import annotations

import ctypes
import os

import torch
import torch.nn as nn

from . import _lib

_ENGINES = {'fp32': _lib.ENGINE_FP32, 'bf16': _lib.ENGINE_BF16, 'bf16x3': _lib.ENGINE_BF16X3}


def _round_up(x, m):
    return (x + m - 1) // m * m


class _Workspace:
    """Zero-initialised scratch per (device, stream); kernels leave it zeroed (include/simclr_b200.h)."""
    _cache = {}

    @classmethod
    def get(cls, device, nbytes, stream=None):
        key = (device.index, stream if stream is not None else _stream(device))
        buf = cls._cache.get(key)
        if buf is None or buf.numel() < nbytes:
            buf = torch.zeros(_round_up(max(nbytes, 1 << 16), 256), dtype=torch.uint8, device=device)
            cls._cache[key] = buf
        return buf


def tc_engine_supported(dim, mask_kind, n_k):
    """Shapes the tcgen05 engine takes; everything else runs on the fp32 engine."""
    return mask_kind != _lib.MASK_EXPLICIT and dim <= 128 and n_k >= 1


def select_engine(requested, feat_dtype, dim, mask_kind, n_k, allow_x3=True):
    req = os.environ.get('SIMCLR_B200_LOSS_ENGINE', requested)
    if req == 'auto':
        if not tc_engine_supported(dim, mask_kind, n_k):
            return _lib.ENGINE_FP32
        # 16-bit features: bf16 operands.  fp32 features (what main_supcon.py:283-290 passes) keep fp32-class
        # arithmetic on the tensor cores: operands split into bf16 hi + lo, three MMAs per tile (1e-4 of the reference)
        if feat_dtype in (torch.bfloat16, torch.float16):
            return _lib.ENGINE_BF16
        return _lib.ENGINE_BF16X3 if allow_x3 else _lib.ENGINE_FP32
    if req not in _ENGINES:
        raise ValueError('Unknown engine: {}'.format(req))
    eng = _ENGINES[req]
    if eng != _lib.ENGINE_FP32 and not tc_engine_supported(dim, mask_kind, n_k):
        raise ValueError('engine {!r} does not take this problem (dim={}, explicit mask={}); '
                         'use engine="fp32" or "auto"'.format(req, dim, mask_kind == _lib.MASK_EXPLICIT))
    return eng


class LossPlan:
    """Shape bookkeeping for one call: where the local rows sit in the (possibly gathered)
    contrast matrix.  Row order is documented in include/simclr_b200.h."""

    def __init__(self, n_local, n_views, dim, n_seg=1, seg_index=0, mode=_lib.MODE_ALL,
                 mask_kind=_lib.MASK_SIMCLR, engine=_lib.ENGINE_FP32):
        self.n_local, self.n_views, self.dim = n_local, n_views, dim
        self.n_seg, self.seg_index = n_seg, seg_index
        self.mode, self.mask_kind, self.engine = mode, mask_kind, engine
        self.rows_local = n_views * n_local
        self.n_k = n_seg * self.rows_local
        self.row_begin = seg_index * self.rows_local
        self.n_anchor_total = self.n_k if mode == _lib.MODE_ALL else n_seg * n_local
        self.scalar_dtype = _lib.F32          # dtype of the loss / upstream-gradient scalars (set by the caller)
        self.row_layout = _lib.ROWS_VIEW_MAJOR
        if engine == _lib.ENGINE_FP32:
            self.k_dtype, self.k_torch, self.ldk = _lib.F32, torch.float32, _round_up(dim, 4)
        else:
            self.k_dtype, self.k_torch, self.ldk = _lib.BF16, torch.bfloat16, _round_up(dim, 64)


def label_ids(labels, device):
    """Labels as int64 class ids whose equality is the reference's `torch.eq(labels, labels.T)` (losses.py:48):
    integer and bool labels are widened, floating-point labels keep their exact value (bit pattern of the
    float64, -0.0 folded onto +0.0) instead of being truncated."""
    labels = labels.to(device=device)
    if labels.is_floating_point():
        return (labels.to(torch.float64) + 0.0).contiguous().view(torch.int64)
    return labels.to(torch.int64).contiguous()


def label_cache_for(lib, prob, device):
    """Device buffer for the label bookkeeping of one forward/backward pair (include/simclr_b200.h), or None."""
    need = lib.scb_supcon_label_cache_bytes(ctypes.byref(prob))
    if need <= 0:
        return None
    cache = torch.empty(need, dtype=torch.uint8, device=device)
    prob.label_cache = cache.data_ptr()
    return cache


def _problem(plan, k, k_lo, labels, mask, temperature, base_temperature):
    p = _lib.SupconProblem()
    p.label_cache = None
    p.k = k.data_ptr()
    p.k_lo = k_lo.data_ptr() if k_lo is not None else None
    p.labels = labels.data_ptr() if labels is not None else None
    p.mask = mask.data_ptr() if mask is not None else None
    p.k_dtype, p.n_k, p.dim, p.ldk = plan.k_dtype, plan.n_k, plan.dim, plan.ldk
    p.n_views, p.seg, p.n_seg = plan.n_views, plan.n_local, plan.n_seg
    p.mask_kind, p.mode, p.engine = plan.mask_kind, plan.mode, plan.engine
    p.scalar_dtype = plan.scalar_dtype
    p.row_layout = plan.row_layout
    p.inv_temperature = 1.0 / temperature
    p.loss_scale = temperature / base_temperature
    return p


def _stream(device):
    # raw handle of torch's current stream (torch.cuda.current_stream(...).cuda_stream builds a Stream object: ~9 us)
    return torch._C._cuda_getCurrentRawStream(device.index if device.index is not None else torch.cuda.current_device())


class _on_device:
    """`with torch.cuda.device(dev)` that costs nothing when `dev` is already current (the usual case:
    one process per GPU); the eager module API is host-bound, every microsecond of Python counts."""
    __slots__ = ('ctx',)

    def __init__(self, dev):
        self.ctx = None if torch.cuda.current_device() == dev.index else torch.cuda.device(dev)

    def __enter__(self):
        if self.ctx is not None:
            self.ctx.__enter__()

    def __exit__(self, *exc):
        if self.ctx is not None:
            self.ctx.__exit__(*exc)
        return False


_WS_BYTES = {}


def _workspace_need(lib, prob, plan):
    """scb_supcon_workspace_bytes, remembered per problem shape."""
    key = (plan.n_k, plan.rows_local, plan.ldk, plan.engine, plan.mask_kind)
    need = _WS_BYTES.get(key)
    if need is None:
        need = _WS_BYTES[key] = lib.scb_supcon_workspace_bytes(ctypes.byref(prob), plan.rows_local)
    return need


def launch_prep(lib, feats, plan, normalize, k_rows, k_lo_rows, inv_norm):
    """feats [n_local, V, D] (last dim contiguous) -> rows of the contrast matrix (scb_supcon_prep)."""
    _lib.check(lib.scb_supcon_prep(
        feats.data_ptr(), _lib.dtype_code(feats.dtype), plan.n_local, plan.n_views, plan.dim,
        feats.stride(0), feats.stride(1), int(normalize), 1e-12,
        k_rows.data_ptr(), k_lo_rows.data_ptr() if k_lo_rows is not None else None,
        plan.k_dtype, plan.ldk, inv_norm.data_ptr() if inv_norm is not None else None,
        _stream(feats.device)), 'scb_supcon_prep')


def launch_forward(lib, prob, plan, lse, possum, npos, loss, device):
    need = _workspace_need(lib, prob, plan)
    st = _stream(device)
    ws = _Workspace.get(device, need, st)
    _lib.check(lib.scb_supcon_forward(
        ctypes.byref(prob), plan.row_begin, plan.rows_local, plan.n_anchor_total,
        lse.data_ptr(), possum.data_ptr(), npos.data_ptr(), loss.data_ptr(),
        ws.data_ptr(), ws.numel(), st), 'scb_supcon_forward')


def launch_backward(lib, prob, plan, lse_all, npos_all, grad_out, dk, device):
    need = _workspace_need(lib, prob, plan)
    ws = _Workspace.get(device, need)
    _lib.check(lib.scb_supcon_backward(
        ctypes.byref(prob), plan.row_begin, plan.rows_local, lse_all.data_ptr(), npos_all.data_ptr(),
        grad_out.data_ptr(), plan.n_anchor_total, dk.data_ptr(), dk.stride(0),
        ws.data_ptr(), ws.numel(), _stream(device)), 'scb_supcon_backward')


def launch_backward_features(lib, prob, plan, lse_all, npos_all, grad_out, feats, normalize, inv_norm, dfeats, device):
    """Backward + finish in one call (tensor-core engines): scb_supcon_backward_features."""
    need = _workspace_need(lib, prob, plan)
    st = _stream(device)
    ws = _Workspace.get(device, need, st)
    _lib.check(lib.scb_supcon_backward_features(
        ctypes.byref(prob), plan.row_begin, plan.rows_local, lse_all.data_ptr(), npos_all.data_ptr(),
        grad_out.data_ptr(), plan.n_anchor_total, feats.data_ptr(), _lib.dtype_code(feats.dtype), plan.n_local,
        feats.stride(0), feats.stride(1), int(normalize), inv_norm.data_ptr() if inv_norm is not None else None,
        dfeats.data_ptr(), dfeats.stride(0), dfeats.stride(1), ws.data_ptr(), ws.numel(), st),
        'scb_supcon_backward_features')


def launch_finish(lib, dk, feats, plan, normalize, inv_norm, dfeats):
    _lib.check(lib.scb_supcon_finish(
        dk.data_ptr(), dk.stride(0), feats.data_ptr(), _lib.dtype_code(feats.dtype),
        plan.n_local, plan.n_views, plan.dim, feats.stride(0), feats.stride(1), int(normalize),
        inv_norm.data_ptr() if inv_norm is not None else None,
        dfeats.data_ptr(), dfeats.stride(0), dfeats.stride(1), _stream(feats.device)),
        'scb_supcon_finish')


def typed_loss(loss2, dtype):
    """The 0-dim loss in the features' dtype without a conversion kernel: the forward leaves the value
    rounded to `scalar_dtype` in the low-order bytes of loss2[1] (include/simclr_b200.h)."""
    if dtype == torch.float32:
        return loss2[0]
    return loss2[1:2].view(dtype)[0]


def grad_scalar(grad_out, dtype):
    """Upstream gradient as a device scalar of the dtype the kernels were told to expect."""
    g = grad_out.detach()
    if g.dtype != dtype:
        g = g.to(dtype)
    return g.contiguous()


class _SupConFunction(torch.autograd.Function):
    """Single-GPU loss: prep -> forward -> (backward -> finish).  losses.py:52-91 + autograd."""

    @staticmethod
    def forward(ctx, feats, labels, mask, plan, temperature, base_temperature, normalize, norms=None):
        lib = _lib.load()
        dev = feats.device
        with _on_device(dev):
            n_k, ldk = plan.n_k, plan.ldk
            # bf16 features that are already laid out as [bsz * n_views, ldk] need neither the stacking copy
            # (losses.py:53) nor a cast: the kernels take the rows in the samples' own order
            zero_copy = (plan.engine == _lib.ENGINE_BF16 and feats.dtype == torch.bfloat16 and not normalize and
                         plan.mode == _lib.MODE_ALL and plan.n_seg == 1 and feats.is_contiguous() and
                         plan.dim == ldk and feats.data_ptr() % 16 == 0)
            plan.row_layout = _lib.ROWS_SAMPLE_MAJOR if zero_copy else _lib.ROWS_VIEW_MAJOR
            k = None if zero_copy else torch.empty((n_k, ldk), dtype=plan.k_torch, device=dev)
            k_lo = torch.empty_like(k) if plan.engine == _lib.ENGINE_BF16X3 else None
            stats = torch.empty((4, n_k), dtype=torch.float32, device=dev)   # lse, possum, npos, inv_norm
            loss = torch.empty(2, dtype=torch.float32, device=dev)     # [fp32 value, value in feats.dtype]
            inv_norm = stats[3] if normalize else None
            if zero_copy:
                k = feats.view(n_k, ldk)            # the feature tensor IS the contrast matrix (sample-major rows)
            else:
                launch_prep(lib, feats, plan, normalize, k, k_lo, inv_norm)
            prob = _problem(plan, k, k_lo, labels, mask, temperature, base_temperature)
            cache = label_cache_for(lib, prob, dev)      # labels are scanned once per step, by the forward
            launch_forward(lib, prob, plan, stats[0], stats[1], stats[2], loss, dev)
            ctx.norms = None
            out = typed_loss(loss, feats.dtype)
            if norms is not None:
                # feature-norm side outputs (main_supcon.py:298-317) from the 1/||z|| vector of the fused normalise
                record, momentum, w_sec, w_l2, terms = norms
                _lib.check(lib.scb_feature_norm_terms(stats[3].data_ptr(), n_k, record.data_ptr(), momentum, w_sec,
                                                      w_l2, terms.data_ptr(), _stream(dev)), 'scb_feature_norm_terms')
                if w_sec != 0.0 or w_l2 != 0.0:
                    out = out + terms[4].to(out.dtype)
                    ctx.norms = (record, w_sec, w_l2)
        ctx.plan, ctx.prob_args, ctx.prob = plan, (temperature, base_temperature, normalize), prob
        ctx.save_for_backward(feats, labels, mask, k, k_lo, stats, cache)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        feats, labels, mask, k, k_lo, stats, _cache = ctx.saved_tensors
        plan = ctx.plan
        temperature, base_temperature, normalize = ctx.prob_args
        lib = _lib.load()
        dev = feats.device
        with _on_device(dev):
            g = grad_scalar(grad_out, feats.dtype)
            prob = ctx.prob          # the struct built by forward (its pointers are kept alive by the saved tensors)
            dfeats = torch.empty(feats.shape, dtype=feats.dtype, device=dev)
            inv_norm = stats[3] if normalize else None
            if plan.engine != _lib.ENGINE_FP32:
                launch_backward_features(lib, prob, plan, stats[0], stats[2], g, feats, normalize, inv_norm, dfeats, dev)
            else:
                dk = torch.empty((plan.rows_local, plan.ldk), dtype=torch.float32, device=dev)
                launch_backward(lib, prob, plan, stats[0], stats[2], g, dk, dev)
                launch_finish(lib, dk, feats, plan, normalize, inv_norm, dfeats)
            if ctx.norms is not None:
                record, w_sec, w_l2 = ctx.norms
                _lib.check(lib.scb_feature_norm_backward(
                    feats.data_ptr(), _lib.dtype_code(feats.dtype), plan.n_local, plan.n_views, plan.dim,
                    feats.stride(0), feats.stride(1), stats[3].data_ptr(), record.data_ptr(), w_sec, w_l2,
                    g.data_ptr(), _lib.dtype_code(g.dtype), dfeats.data_ptr(), dfeats.stride(0), dfeats.stride(1),
                    _stream(dev)), 'scb_feature_norm_backward')
        return dfeats, None, None, None, None, None, None, None


class _GraphEntry:
    """One captured forward+backward of the loss for a fixed (shape, dtype, options, input address)."""
    __slots__ = ('graph', 'x', 'labels', 'ones', 'loss', 'grad', 'serial', 'own_input')

    def launch(self, feats):
        """Issues the captured step for `feats` (copied into the owned buffer unless the graph is bound to its address)."""
        if self.own_input:
            self.x.copy_(feats.detach())
        self.graph.replay()
        self.serial += 1


class _GraphedSupCon(torch.autograd.Function):
    """Replays a CUDA graph that holds the forward AND the backward kernels of one loss call (upstream gradient 1);
    autograd's backward then only scales the stored gradient.  Cuts the host side of a step from four C-ABI calls
    plus autograd bookkeeping to one graph launch."""

    @staticmethod
    def forward(ctx, feats, entry, launched=False):
        if not launched:
            entry.launch(feats)
        ctx.entry, ctx.serial = entry, entry.serial
        # the graph's own output buffer (as torch.cuda.make_graphed_callables does): no copy kernel; the next call of
        # the same shape overwrites it
        return entry.loss.detach()

    @staticmethod
    def backward(ctx, grad_out):
        e = ctx.entry
        if e.serial != ctx.serial:
            raise RuntimeError('SupConLoss(graph=True): backward of a forward that was followed by another forward of '
                               'the same shape (the captured gradient buffer was overwritten); call backward first '
                               'or use graph=False')
        return e.grad * grad_out.to(e.grad.dtype), None, None


def validate_inputs(features, labels, mask, contrast_mode):
    """The reference's argument checks, in its order, with its messages (losses.py:33-61)."""
    if len(features.shape) < 3:
        raise ValueError('`features` needs to be [bsz, n_views, ...],'
                         'at least 3 dimensions are required')
    batch_size = features.shape[0]
    if labels is not None and mask is not None:
        raise ValueError('Cannot define both `labels` and `mask`')
    if labels is not None:
        labels = labels.contiguous().view(-1)
        if labels.shape[0] != batch_size:
            raise ValueError('Num of labels does not match num of features')
    if contrast_mode == 'one':
        mode = _lib.MODE_ONE
    elif contrast_mode == 'all':
        mode = _lib.MODE_ALL
    else:
        raise ValueError('Unknown mode: {}'.format(contrast_mode))
    return labels, mode


class SupConLoss(nn.Module):
    """Supervised contrastive loss; SimCLR's NT-Xent when `labels` and `mask` are both None.

    Mirrors /root/reference/losses.py:7-93 (constructor :10-15, forward :17-93).
    """

    def __init__(self, temperature=0.07, contrast_mode='all', base_temperature=0.07,
                 engine='auto', normalize=False, track_norms=False, norm_momentum=0.1, graph=False):
        super(SupConLoss, self).__init__()
        # graph=True: forward + backward of a call are captured once per (shape, dtype, labels?) in a CUDA graph and
        # replayed afterwards (see _GraphedSupCon); for training loops whose step is bound by host launch overhead.
        # As with torch.cuda.make_graphed_callables the returned loss is the graph's own output buffer: read it
        # (.item(), a copy) before the next call with the same shape overwrites it.
        self.graph = graph
        self._graph_cache, self._graph_cache_settings, self._graph_hot = None, None, {}
        self.temperature = temperature
        self.contrast_mode = contrast_mode
        self.base_temperature = base_temperature
        self.engine = engine
        self.normalize = normalize
        # feature-norm side outputs of the training loop (main_supcon.py:298-317); need normalize=True (raw features in)
        self.track_norms = track_norms
        self.norm_momentum = norm_momentum           # `--norm_momentum` of the reference
        self.record_norm_mean = None                 # device [value, initialised flag]: `opt.record_norm_mean`
        self.norm_terms = None                       # device [norm_mean, norm_var, loss_sec, loss_l2reg, added term]

    def forward(self, features, labels=None, mask=None, sec_weight=0.0, l2reg_weight=0.0):
        """features [bsz, n_views, ...]; labels [bsz]; mask [bsz, bsz] (float, may be asymmetric).
        Returns the 0-dim loss, differentiable with respect to `features`.

        With `track_norms=True` (and `normalize=True`) the call also leaves the feature-norm statistics of
        main_supcon.py:298-317 in `self.norm_terms` (a device tensor, no synchronisation), keeps the moving average in
        `self.record_norm_mean`, and adds `sec_weight * loss_sec + l2reg_weight * loss_l2reg` to the loss (the
        reference's `opt.sec_wei * now_iter / total` and `opt.l2reg_wei * ...` factors, both 0 by default)."""
        if self.graph and labels is None and mask is None and sec_weight == 0.0 and l2reg_weight == 0.0:
            # graph mode, recurring input buffer: one dictionary probe, then the replay (everything below ran when the
            # graph for this buffer was captured; the probe compares all it depends on)
            hot = self._graph_hot.get(features.data_ptr())
            if (hot is not None and features.shape == hot[0] and features.stride() == hot[1] and
                    features.dtype is hot[2] and self._graph_settings() == hot[3] and features.requires_grad and
                    torch.is_grad_enabled() and not torch.cuda.is_current_stream_capturing()):
                hot[4].launch(features)         # the GPU starts now; autograd's bookkeeping for the call runs beside it
                return _GraphedSupCon.apply(features, hot[4], True)
        labels, mode = validate_inputs(features, labels, mask, self.contrast_mode)
        if not features.is_cuda:
            raise RuntimeError('simclr_pytorch_distributed_b200.SupConLoss runs on CUDA tensors only '
                               '(there is no CPU fallback); got a %s tensor' % features.device.type)
        if len(features.shape) > 3:
            features = features.view(features.shape[0], features.shape[1], -1)     # losses.py:36-37
        own_buffer = True
        if features.stride(2) != 1 and features.shape[2] > 1:
            features, own_buffer = features.contiguous(), False
        B, V, D = features.shape
        dev = features.device
        if not self.track_norms and (sec_weight != 0.0 or l2reg_weight != 0.0):
            raise ValueError('sec_weight / l2reg_weight need track_norms=True')
        if (self.graph and mask is None and not self.track_norms and features.requires_grad and
                torch.is_grad_enabled() and not torch.cuda.is_current_stream_capturing()):
            # graph mode: nothing below is needed per call (the captured step was planned once)
            return self._graphed(features, label_ids(labels, dev) if labels is not None else None, own_buffer)
        if labels is not None:
            mask_kind = _lib.MASK_LABELS
            labels = label_ids(labels, dev)
        elif mask is not None:
            mask_kind = _lib.MASK_EXPLICIT
            mask = mask.to(device=dev, dtype=torch.float32).contiguous()            # losses.py:50
            if mask.shape != (B, B):
                raise ValueError('`mask` must be [bsz, bsz]')
        else:
            mask_kind = _lib.MASK_SIMCLR
        engine = select_engine(self.engine, features.dtype, D, mask_kind, V * B)
        plan = LossPlan(B, V, D, mode=mode, mask_kind=mask_kind, engine=engine)
        plan.scalar_dtype = _lib.dtype_code(features.dtype)
        norms = None
        if self.track_norms:
            if not self.normalize:
                raise ValueError('track_norms needs normalize=True: the norms are those of the un-normalised features')
            if self.record_norm_mean is None or self.record_norm_mean.device != dev:
                self.record_norm_mean = torch.zeros(2, dtype=torch.float32, device=dev)
            self.norm_terms = torch.empty(5, dtype=torch.float32, device=dev)
            norms = (self.record_norm_mean, float(self.norm_momentum), float(sec_weight), float(l2reg_weight),
                     self.norm_terms)
        return _SupConFunction.apply(features, labels, mask, plan, float(self.temperature),
                                     float(self.base_temperature), bool(self.normalize), norms)

    # ------------------------------------------------------------------ graph mode
    def _graph_settings(self):
        """What a captured graph has baked in besides the input's shape: changing any of it drops the graphs."""
        return (self.temperature, self.contrast_mode, self.base_temperature, self.engine, self.normalize, self.track_norms)

    def _graphed(self, features, labels, bind=True):
        settings = self._graph_settings()
        if self._graph_cache is None or self._graph_cache_settings != settings:
            self._graph_cache = GraphCache(lambda: SupConLoss(self.temperature, self.contrast_mode, self.base_temperature,
                                                              self.engine, self.normalize))
            self._graph_cache_settings = settings
            self._graph_hot = {}
        out = self._graph_cache.run(features, labels, bind)
        entry = self._graph_cache.last_entry
        if labels is None and not entry.own_input:          # bound to this buffer: later calls take the probe in forward
            self._graph_hot[features.data_ptr()] = (features.shape, features.stride(), features.dtype, settings, entry)
        return out


class GraphCache:
    """Captured forward+backward graphs of a loss module, per (shape, dtype, strides, labels?) and input address.

    A graph is bound to the address of its input: callers that feed the loss from a few recurring addresses (a
    double-buffered prefetcher, a preallocated projection output, or simply a training loop whose features land at the
    same allocator address every step) get one graph per address and no copy; any other input is copied into a buffer
    the cache owns.  `bind_addresses=False` (the row-sharded loss): always the owned
    buffer, so that whether a call captures or replays depends on the shape alone and every rank decides alike."""
    MAX_ADDRESSES = 4           # captured graphs bound to the caller's own input buffers, per shape

    def __init__(self, make_inner, bind_addresses=True):
        self.make_inner = make_inner
        self.bind_addresses = bind_addresses
        self.graphs = {}
        self.last_entry = None      # the entry the last run() used

    def run(self, features, labels, bind=True):
        """`bind=False`: this input is a temporary of the caller (its address will not recur): use the owned buffer."""
        dev = features.device
        base = (tuple(features.shape), features.dtype, dev.index, tuple(features.stride()), labels is not None)
        per_shape = self.graphs.setdefault(base, {})
        addr = features.data_ptr()
        bind = bind and self.bind_addresses
        entry = per_shape.get(addr) if bind else None
        if entry is None:
            bound = [k for k in per_shape if k != 'own']
            if bind and len(bound) < self.MAX_ADDRESSES and features.is_contiguous():
                entry = self._capture(features, labels, own_input=False)
                per_shape[addr] = entry
            else:
                entry = per_shape.get('own')
                if entry is None:
                    entry = per_shape['own'] = self._capture(features, labels, own_input=True)
        if labels is not None:
            entry.labels.copy_(labels)
        self.last_entry = entry
        return _GraphedSupCon.apply(features, entry)

    def _capture(self, features, labels, own_input):
        inner = self.make_inner()
        e = _GraphEntry()
        e.own_input, e.serial = own_input, 0
        e.x = (torch.empty_like(features) if own_input else features.detach())
        e.labels = labels.clone() if labels is not None else None
        leaf = e.x.detach().requires_grad_(True)
        if own_input:
            e.x.copy_(features.detach())
        # every tensor the graph reads by address lives as long as the entry: input, labels, the upstream-gradient scalar
        ones = e.ones = torch.ones((), dtype=features.dtype, device=features.device)
        side = torch.cuda.Stream(device=features.device)
        side.wait_stream(torch.cuda.current_stream(features.device))
        with torch.cuda.stream(side):                      # warm-up outside the capture: builds, workspaces, descriptors
            for _ in range(2):
                leaf.grad = None
                inner(leaf, labels=e.labels).backward(ones)
        torch.cuda.current_stream(features.device).wait_stream(side)
        leaf.grad = None
        e.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(e.graph):
            loss = inner(leaf, labels=e.labels)
            loss.backward(ones)
        e.loss, e.grad = loss.detach(), leaf.grad
        if not own_input:
            # Do not keep the caller's buffer alive: a training loop's features are a fresh allocation every step that
            # the caching allocator places at the same address once the previous step's tensor is gone -- holding on to
            # it would force a new address (and a new capture) every step.  The graph is only ever replayed for an
            # incoming tensor that IS at this address (same shape, strides, dtype), so the memory is live when it runs.
            e.x = None
        return e
