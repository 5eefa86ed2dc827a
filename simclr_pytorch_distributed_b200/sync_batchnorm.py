"""`SyncBatchNorm` -- drop-in for the `torch.nn.SyncBatchNorm` layers that the reference's
`--syncBN` flag swaps into the network (/root/reference/main_supcon.py:69-70,222-224; BN sites
/root/reference/networks/resnet_big.py:14,16,22,44,46,48,54,77).

Same constructor, parameters, buffers and state-dict keys as `torch.nn.SyncBatchNorm`
(torch/nn/modules/batchnorm.py:718-733) and the same `convert_sync_batchnorm` classmethod
(:844-905); the arithmetic (torch/nn/modules/_functions.py:10-209) runs in the sm_100a kernels of
csrc/bn.cu and csrc/bn_cluster.cu through the C ABI.  Per layer and pass: one launch on one rank when the
tensor fits the registers of a resident wave, else one streaming reduction, one tiny peer-memory
exchange (only when world_size > 1) and one fused apply -- and no host synchronisation (stock
SyncBatchNorm forces one per layer, _functions.py:88-100).

CUDA tensors only; there is no CPU path.  `fuse_relu=True` additionally applies ReLU inside the
apply kernels (the `relu(bn(x))` sites of resnet_big.py:26,58,59,111), and
`forward(x, z)` adds the residual operand of resnet_big.py:27-30,60-63 before that ReLU.
"""
from __future__ import annotations

import torch
import torch.distributed as dist
from torch.nn.modules.batchnorm import _BatchNorm

from . import _lib


class _Kernels:
    """Thin typed wrappers over the C ABI (one per entry point of include/simclr_b200.h)."""

    _ws = {}
    _ws_need = {}
    _fused_cache = {}

    @classmethod
    def workspace(cls, lib, device, channels):
        need = cls._ws_need.get(channels)
        if need is None:
            need = cls._ws_need[channels] = lib.scb_bn_workspace_bytes(channels)
        key = (device.index, cls.stream(device))
        buf = cls._ws.get(key)
        if buf is None or buf.numel() < need:
            buf = torch.zeros(max(need, 1 << 16), dtype=torch.uint8, device=device)
            cls._ws[key] = buf
        return buf

    @staticmethod
    def stream(device):
        # the raw handle of torch's current stream: torch.cuda.current_stream(device).cuda_stream costs ~9 us a call
        # (it builds a Stream object), and a forward+backward asks five times
        return torch._C._cuda_getCurrentRawStream(device.index if device.index is not None else torch.cuda.current_device())

    @classmethod
    def fused_supported(cls, x, geom, which, flags=0):
        """True when this tensor fits the single-launch kernels (which: 0 forward, 1 backward)."""
        import os
        if os.environ.get('SCB_BN_FUSED', '1') == '0':          # tests / measurements: force the multi-kernel path
            return False
        key = (x.dtype, geom, which, flags, x.device.index)
        hit = cls._fused_cache.get(key)
        if hit is None:
            lib = _lib.load()
            cl, outer, C, inner = geom
            with torch.cuda.device(x.device):
                hit = bool(lib.scb_bn_fused_supported(_lib.dtype_code(x.dtype), cl, outer, C, inner, which, flags))
            cls._fused_cache[key] = hit
        return hit

    @staticmethod
    def param_dtype(*tensors):
        """dtype code of the per-channel parameter tensors of a call (weight, bias, running statistics, parameter
        gradients): they must agree; fp32 when there are none."""
        code = None
        for t in tensors:
            if t is None:
                continue
            c = _lib.dtype_code(t.dtype)
            if code is None:
                code = c
            elif c != code:
                raise ValueError('SyncBatchNorm kernels need weight, bias, running statistics and their gradients in one '
                                 'dtype (got %s)' % sorted({str(u.dtype) for u in tensors if u is not None}))
        return _lib.F32 if code is None else code

    @classmethod
    def stats(cls, x, geom, out_stats):
        lib = _lib.load()
        cl, outer, C, inner = geom
        if cls.fused_supported(x, geom, 0) and x.data_ptr() % 16 == 0:      # cluster reduction, no scratch
            rc = lib.scb_bn_stats_fused(x.data_ptr(), _lib.dtype_code(x.dtype), outer * inner, C,
                                        out_stats.data_ptr(), cls.stream(x.device))
            if rc == 0:
                return
            if rc != -3:
                _lib.check(rc, 'scb_bn_stats_fused')
        ws = cls.workspace(lib, x.device, C)
        _lib.check(lib.scb_bn_stats(x.data_ptr(), _lib.dtype_code(x.dtype), cl, outer, C, inner,
                                    out_stats.data_ptr(), ws.data_ptr(), ws.numel(),
                                    cls.stream(x.device)), 'scb_bn_stats')

    @classmethod
    def apply(cls, x, y, geom, stats_all, n_ranks, weight, bias, running_mean, running_var,
              momentum, eps, saved, flags, z=None, nbt=None):
        lib = _lib.load()
        cl, outer, C, inner = geom
        p = lambda t: t.data_ptr() if t is not None else None   # noqa: E731
        # saved = [mean (C) | invstd (C) | count]: plain pointer arithmetic (a slice is a new tensor object, ~2 us)
        s0 = saved.data_ptr() if saved is not None else None
        sm, si, sc = (s0, s0 + 4 * C, s0 + 8 * C) if s0 is not None else (None, None, None)
        _lib.check(lib.scb_bn_apply(x.data_ptr(), p(z), y.data_ptr(), _lib.dtype_code(x.dtype), cl, outer, C,
                                    inner, p(stats_all), n_ranks, p(weight), p(bias), p(running_mean),
                                    p(running_var), momentum, eps, sm, si, sc, p(nbt), flags,
                                    cls.param_dtype(weight, bias, running_mean, running_var),
                                    cls.stream(x.device)), 'scb_bn_apply')

    @classmethod
    def backward_reduce(cls, dy, x, geom, saved, weight, bias, sums, grad_weight, grad_bias, flags, yout=None):
        lib = _lib.load()
        cl, outer, C, inner = geom
        p = lambda t: t.data_ptr() if t is not None else None   # noqa: E731
        if (cls.fused_supported(x, geom, 1, flags) and x.data_ptr() % 16 == 0 and dy.data_ptr() % 16 == 0 and
                (yout is None or yout.data_ptr() % 16 == 0)):
            rc = lib.scb_bn_backward_reduce_fused(
                dy.data_ptr(), x.data_ptr(), p(yout), _lib.dtype_code(x.dtype), outer * inner, C,
                saved.data_ptr(), saved.data_ptr() + 4 * C, p(weight), p(bias), saved.data_ptr() + 8 * C,
                sums.data_ptr(), p(grad_weight), p(grad_bias), flags,
                cls.param_dtype(weight, bias, grad_weight, grad_bias), cls.stream(x.device))
            if rc == 0:
                return
            if rc != -3:
                _lib.check(rc, 'scb_bn_backward_reduce_fused')
        ws = cls.workspace(lib, x.device, C)
        _lib.check(lib.scb_bn_backward_reduce(
            dy.data_ptr(), x.data_ptr(), p(yout), _lib.dtype_code(x.dtype), cl, outer, C, inner,
            saved.data_ptr(), saved.data_ptr() + 4 * C, p(weight), p(bias), sums.data_ptr(),
            p(grad_weight), p(grad_bias), flags, cls.param_dtype(weight, bias, grad_weight, grad_bias),
            ws.data_ptr(), ws.numel(), cls.stream(x.device)), 'scb_bn_backward_reduce')

    @classmethod
    def forward_fused(cls, x, y, geom, weight, bias, running_mean, running_var, momentum, eps, saved, flags, z=None,
                      nbt=None):
        lib = _lib.load()
        cl, outer, C, inner = geom
        p = lambda t: t.data_ptr() if t is not None else None   # noqa: E731
        ws = cls.workspace(lib, x.device, C)
        return lib.scb_bn_forward_fused(x.data_ptr(), p(z), y.data_ptr(), _lib.dtype_code(x.dtype), outer * inner, C,
                                        p(weight), p(bias), p(running_mean), p(running_var), momentum, eps,
                                        saved.data_ptr(), saved.data_ptr() + 4 * C, saved.data_ptr() + 8 * C,
                                        p(nbt), flags, cls.param_dtype(weight, bias, running_mean, running_var),
                                        ws.data_ptr(), ws.numel(), cls.stream(x.device))

    @classmethod
    def backward_fused(cls, dy, x, dx, geom, saved, weight, bias, sums, grad_weight, grad_bias, flags,
                       yout=None, dz=None):
        lib = _lib.load()
        cl, outer, C, inner = geom
        p = lambda t: t.data_ptr() if t is not None else None   # noqa: E731
        ws = cls.workspace(lib, x.device, C)
        return lib.scb_bn_backward_fused(dy.data_ptr(), x.data_ptr(), p(yout), dx.data_ptr(), p(dz),
                                         _lib.dtype_code(x.dtype), outer * inner, C, saved.data_ptr(),
                                         saved.data_ptr() + 4 * C, p(weight), p(bias), saved.data_ptr() + 8 * C,
                                         sums.data_ptr(), p(grad_weight), p(grad_bias), flags,
                                         cls.param_dtype(weight, bias, grad_weight, grad_bias), ws.data_ptr(),
                                         ws.numel(), cls.stream(x.device))

    @classmethod
    def backward_apply(cls, dy, x, dx, geom, saved, weight, bias, sums, flags, yout=None, dz=None):
        lib = _lib.load()
        cl, outer, C, inner = geom
        p = lambda t: t.data_ptr() if t is not None else None   # noqa: E731
        _lib.check(lib.scb_bn_backward_apply(
            dy.data_ptr(), x.data_ptr(), p(yout), dx.data_ptr(), p(dz), _lib.dtype_code(x.dtype), cl, outer, C, inner,
            saved.data_ptr(), saved.data_ptr() + 4 * C, p(weight), p(bias), sums.data_ptr(),
            saved.data_ptr() + 8 * C, flags, cls.param_dtype(weight, bias), cls.stream(x.device)),
            'scb_bn_backward_apply')


def _is_channels_last(x):
    return ((x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last)) or
            (x.dim() == 5 and x.is_contiguous(memory_format=torch.channels_last_3d)))


def _layout(x):
    """Return (tensor in a supported layout, (channels_last, outer, C, inner)) -- the same
    "keep channels_last, else make contiguous" rule as _functions.py:22-26."""
    C = x.size(1)
    if x.dim() == 2:
        x = x.contiguous()
        return x, (1, x.size(0), C, 1)                 # [N, C] is already "rows x channels"
    if _is_channels_last(x):
        return x, (1, x.numel() // C if C else 0, C, 1)
    x = x.contiguous()
    inner = x.numel() // (x.size(0) * C) if x.size(0) * C else 1
    return x, (0, x.size(0), C, max(inner, 1))


def _match_layout(t, like, geom):
    """grad_output in the same memory layout as the saved input (_functions.py:129-133)."""
    if geom[0] == 1 and like.dim() == 4:
        return t.contiguous(memory_format=torch.channels_last)
    if geom[0] == 1 and like.dim() == 5:
        return t.contiguous(memory_format=torch.channels_last_3d)
    return t.contiguous()


def _f32(t):
    return t if t is None or t.dtype == torch.float32 else t.float()


class _PeerStats:
    """The per-channel exchanges (_functions.py:41-56 all_gather, :155-165 all_reduce) over peer memory.

    A few kilobytes per layer: an NCCL collective is pure latency here.  Each rank owns symmetric-memory
    buffers (torch.distributed._symmetric_memory, plumbing only); `scb_peer_bcast` stores the rank's message
    into every rank's copy and the handle's stream barrier orders it against the readers.  Buffers are shared
    by all layers with the same message length and rotate through GENERATIONS copies (a rank is never more
    than one exchange ahead of its slowest peer).  SCB_P2P=0 or any set-up failure -> NCCL.

    The kernels of this library never wait on another GPU: the cross-rank wait is the stream barrier between two
    launches, so a concurrent NCCL kernel (DDP's bucket all-reduce) can neither starve nor be starved by them."""
    GENERATIONS = 4
    _cache = {}
    _failed = False

    def __init__(self, group, dev, n, world, rank):
        import ctypes
        import torch.distributed._symmetric_memory as symm
        self.world, self.n = world, n
        self.bufs, self.ptrs, self.handles = [], [], []
        for _ in range(self.GENERATIONS):
            b = symm.empty((world * n,), dtype=torch.float32, device=dev)
            h = symm.rendezvous(b, group)
            self.bufs.append(b)
            self.handles.append(h)
            self.ptrs.append((ctypes.c_void_p * world)(*[int(p) for p in h.buffer_ptrs]))
        self.seg = (ctypes.c_int64 * 3)(0, rank * n * 4, n * 4)
        self.calls = 0
        self.capture_id, self.capture_calls = 0, 0
        self.rank = rank

    @classmethod
    def get(cls, group, dev, n, world, rank):
        import os
        if cls._failed or os.environ.get('SCB_P2P', '1') == '0' or world > 16 or dev.type != 'cuda':
            return None
        if group._get_backend_name() != 'nccl':
            return None
        key = (id(group), dev.index, n, world)
        if key not in cls._cache:
            try:
                cls._cache[key] = cls(group, dev, n, world, rank)
            except Exception as e:       # every rank fails alike -> NCCL everywhere
                import warnings
                cls._failed = True
                warnings.warn('peer-memory exchange unavailable (%s); SyncBatchNorm uses NCCL' % (e,))
                return None
        return cls._cache[key]

    def all_gather(self, src, dev):
        """src: n floats of this rank -> [world * n] floats, rank-major (valid until GENERATIONS-1 further calls)."""
        lib = _lib.load()
        # Buffer rotation: eager calls walk through the generations.  A CUDA-graph capture bakes the choice in, so the
        # rotation restarts at every capture and the first exchange of a graph (and the first eager one after replays)
        # is preceded by a barrier: no replay seam can hand a rank a buffer that a peer is still reading.
        cap = lib.scb_stream_capture_id(_Kernels.stream(dev))
        pre_barrier = False
        if cap > 0:
            if cap != self.capture_id:
                self.capture_id, self.capture_calls, pre_barrier = cap, 0, True
            g = self.capture_calls % self.GENERATIONS
            self.capture_calls += 1
        else:
            if self.capture_id != 0:
                self.capture_id, pre_barrier = 0, True
            g = self.calls % self.GENERATIONS
        self.calls += 1
        if pre_barrier:
            self.handles[g].barrier(channel=0)
        _lib.check(lib.scb_peer_bcast(src.data_ptr(), self.seg, 1, self.ptrs[g], self.world,
                                      _Kernels.stream(dev)), 'scb_peer_bcast')
        self.handles[g].barrier(channel=0)
        return self.bufs[g]


def _exchange(kern, vec, n, process_group, world_size, dev):
    """All-gather of one per-rank message of n floats -> [world_size * n] floats, rank-major."""
    px = (_PeerStats.get(process_group, dev, n, world_size, dist.get_rank(process_group))
          if kern is _Kernels else None)
    if px is not None:
        return px.all_gather(vec, dev)
    out = torch.empty(world_size * n, dtype=torch.float32, device=dev)
    if process_group._get_backend_name() != 'gloo':
        dist.all_gather_into_tensor(out, vec, process_group, async_op=False)
    else:                                            # gloo has no all_gather_into_tensor
        parts = [torch.empty_like(vec) for _ in range(world_size)]
        dist.all_gather(parts, vec, process_group, async_op=False)
        out = torch.cat(parts)
    return out


class _SyncBNFunction(torch.autograd.Function):
    """Training-mode forward/backward: _functions.py:10-209 on the sm_100a kernels.

    `z` (optional) is a residual operand added after the normalisation and before the fused ReLU
    (resnet_big.py:27-30,60-63): y = relu(bn(x) + z)."""

    kernels = _Kernels      # tests swap this for an oracle-backed stand-in to check the host logic

    @staticmethod
    def forward(ctx, input, z, weight, bias, running_mean, running_var, eps, momentum, process_group,
                world_size, fuse_relu, use_batch_stats, nbt=None):
        # nbt: the module's num_batches_tracked buffer when the kernels are to count the step (training, momentum given)
        nbt_kw = {'nbt': nbt} if nbt is not None else {}
        kern = _SyncBNFunction.kernels
        x, geom = _layout(input)
        C = geom[2]
        size = int(x.numel() // C)
        if use_batch_stats and size == 1 and world_size < 2:
            raise ValueError(f"Expected more than 1 value per channel when training, got input size {size}")
        # The kernels take weight, bias and the running statistics in the module's own dtype (fp32, bf16 or fp16) as long
        # as they agree: a model converted with .to(torch.bfloat16) then needs no conversion kernels around the passes.
        # Mixed dtypes (rare) go through fp32 copies.  The oracle-backed stand-in of the CPU tests is fp32-only.
        ptensors = [t for t in (weight, bias, running_mean, running_var) if t is not None]
        native = (kern is _Kernels and len({t.dtype for t in ptensors}) <= 1 and
                  all(t.dtype in (torch.float32, torch.bfloat16, torch.float16) and t.is_contiguous() for t in ptensors))
        if native:
            w32, b32, rm, rv = weight, bias, running_mean, running_var
        else:
            w32, b32 = _f32(weight), _f32(bias)
            rm, rv = _f32(running_mean), _f32(running_var)
        dev = x.device
        flags = (_lib.BN_RELU if fuse_relu else 0) | (_lib.BN_ADD if z is not None else 0)
        if z is not None:
            if z.shape != input.shape:
                raise ValueError('the residual operand must have the shape of the input')
            z = _match_layout(z, x, geom)
            if z.dtype != x.dtype:
                z = z.to(x.dtype)
        y = torch.empty_like(x)
        saved = torch.empty(2 * C + 1, dtype=torch.float32, device=dev)     # mean, invstd, n_total
        done = False
        if (use_batch_stats and world_size == 1 and hasattr(kern, 'forward_fused') and
                kern.fused_supported(x, geom, 0, flags)):
            # one rank, tensor fits one resident wave: statistics + normalisation in a single launch
            rc = kern.forward_fused(x, y, geom, w32, b32, rm, rv, momentum, eps, saved, flags, z=z, **nbt_kw)
            if rc == 0:
                done = True
            elif rc != -3:
                _lib.check(rc, 'scb_bn_forward_fused')
        if not done:
            if not use_batch_stats:
                stats_all = None            # eval: the running statistics are the statistics
            else:
                # message of this rank: [mean(C), M2(C), count]  (_functions.py:41-49 sends mean, invstd, count)
                stats_all = torch.empty(2 * C + 1, dtype=torch.float32, device=dev)
                kern.stats(x, geom, stats_all)
                if world_size > 1:
                    stats_all = _exchange(kern, stats_all, 2 * C + 1, process_group, world_size, dev)
            # BN_PDL: x was complete before the statistics pass started, so the apply pass may start filling its ring
            # while that pass (or the exchange behind it) is still retiring
            if z is not None:
                kern.apply(x, y, geom, stats_all, world_size, w32, b32, rm, rv, momentum, eps, saved, flags, z=z, **nbt_kw)
            else:
                kern.apply(x, y, geom, stats_all, world_size, w32, b32, rm, rv, momentum, eps, saved,
                           flags | (_lib.BN_PDL if use_batch_stats else 0), **nbt_kw)
        if rm is not running_mean and running_mean is not None:
            running_mean.copy_(rm)
        if rv is not running_var and running_var is not None:
            running_var.copy_(rv)
        gate_by_y = fuse_relu and z is not None
        ctx.save_for_backward(x, w32, b32, saved, y if gate_by_y else None)
        ctx.geom, ctx.flags, ctx.use_batch_stats = geom, flags, use_batch_stats
        ctx.has_z = z is not None
        ctx.process_group, ctx.world_size = process_group, world_size
        ctx.param_dtypes = (None if weight is None else weight.dtype, None if bias is None else bias.dtype)
        return y

    @staticmethod
    def backward(ctx, grad_output):
        kern = _SyncBNFunction.kernels
        x, w32, b32, saved, yout = ctx.saved_tensors
        geom, flags = ctx.geom, ctx.flags
        C = geom[2]
        dy = _match_layout(grad_output, x, geom)
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        dev = x.device
        sums = torch.empty(2 * C, dtype=torch.float32, device=dev)          # sum dy, sum dy (x - mean)
        pdt = w32.dtype if w32 is not None else (b32.dtype if b32 is not None else torch.float32)
        gwb = torch.empty(2 * C, dtype=pdt, device=dev)                      # dgamma, dbeta in the parameters' dtype
        gw, gb = gwb[0:C], gwb[C:]
        need_dx = ctx.needs_input_grad[0]
        need_dz = ctx.has_z and ctx.needs_input_grad[1]
        gated = yout is not None                                             # ReLU after the residual add
        grad_input = torch.empty_like(x) if need_dx else None
        grad_z = torch.empty_like(x) if (need_dz and gated) else None
        kw = {}
        if gated:
            kw['yout'] = yout
        done = False
        if (need_dx and ctx.use_batch_stats and ctx.world_size == 1 and hasattr(kern, 'backward_fused') and
                kern.fused_supported(x, geom, 1, flags)):
            rc = kern.backward_fused(dy, x, grad_input, geom, saved, w32, b32, sums, gw, gb, flags, yout=yout,
                                     dz=grad_z)
            if rc == 0:
                done = True
            elif rc != -3:
                _lib.check(rc, 'scb_bn_backward_fused')
        if not done:
            kern.backward_reduce(dy, x, geom, saved, w32, b32, sums, gw, gb, flags, **kw)
            if need_dx or grad_z is not None:
                if not ctx.use_batch_stats:
                    sums = torch.zeros_like(sums)               # eval: statistics do not depend on x
                elif ctx.world_size > 1:                        # _functions.py:155-165: all-reduce = all-gather + sum
                    sums = _exchange(kern, sums, 2 * C, ctx.process_group, ctx.world_size, dev) \
                        .view(ctx.world_size, 2 * C).sum(0)
                if grad_input is None:
                    grad_input = torch.empty_like(x)
                if gated:
                    kw['dz'] = grad_z
                kern.backward_apply(dy, x, grad_input, geom, saved, w32, b32, sums, flags | _lib.BN_PDL, **kw)
        if need_dz and not gated:
            grad_z = dy                                          # plain add: the gradient passes through
        if not need_dx:
            grad_input = None
        wd, bd = ctx.param_dtypes
        # dgamma / dbeta stay un-synced: DDP averages them (_functions.py:180-181)
        grad_weight = gw.to(wd) if (w32 is not None and ctx.needs_input_grad[2]) else None
        grad_bias = gb.to(bd) if (b32 is not None and ctx.needs_input_grad[3]) else None
        return grad_input, grad_z, grad_weight, grad_bias, None, None, None, None, None, None, None, None, None


class SyncBatchNorm(_BatchNorm):
    """Batch normalisation with statistics over the whole (cross-rank) mini-batch.

    Argument meaning, buffers and error behaviour follow `torch.nn.SyncBatchNorm`
    (torch/nn/modules/batchnorm.py:615-905).  Unlike the stock module it runs the same kernels at
    world size 1 instead of falling back to cuDNN (:818-828), so one code path serves 1..8 GPUs.
    """

    def __init__(self, num_features, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True,
                 process_group=None, device=None, dtype=None, fuse_relu=False):
        factory_kwargs = {'device': device, 'dtype': dtype}
        super().__init__(num_features, eps, momentum, affine, track_running_stats, **factory_kwargs)
        self.process_group = process_group
        self.fuse_relu = fuse_relu

    def _check_input_dim(self, input):
        if input.dim() < 2:
            raise ValueError(f"expected at least 2D input (got {input.dim()}D input)")

    def _check_non_zero_input_channels(self, input):
        if input.size(1) == 0:
            raise ValueError("SyncBatchNorm number of input channels should be non-zero")

    def forward(self, input, z=None):
        """`z`: optional residual operand, y = bn(input) + z (then ReLU with fuse_relu=True): the tail of a
        residual block, resnet_big.py:27-30,60-63, in one pass."""
        self._check_input_dim(input)
        self._check_non_zero_input_channels(input)
        if input.size(1) != self.num_features:
            raise ValueError(f"expected {self.num_features} channels, got {input.size(1)}")
        if not input.is_cuda:
            raise ValueError("SyncBatchNorm expected input tensor to be on GPU "
                             "(simclr_pytorch_distributed_b200 has no CPU path)")

        if self.momentum is None:
            exponential_average_factor = 0.0
        else:
            exponential_average_factor = self.momentum
        nbt = None
        if self.training and self.track_running_stats:
            if self.momentum is None:                     # cumulative moving average: the host needs the count
                self.num_batches_tracked.add_(1)
                exponential_average_factor = 1.0 / self.num_batches_tracked.item()
            else:
                exponential_average_factor = self.momentum
                nbt = self.num_batches_tracked            # counted by the normalising kernel: no launch of its own

        if self.training:
            bn_training = True
        else:
            bn_training = (self.running_mean is None) and (self.running_var is None)
        running_mean = self.running_mean if not self.training or self.track_running_stats else None
        running_var = self.running_var if not self.training or self.track_running_stats else None

        world_size, process_group = 1, None
        # statistics are exchanged only in training mode (batchnorm.py:791-815)
        if bn_training and self.training and dist.is_available() and dist.is_initialized():
            process_group = self.process_group if self.process_group else dist.group.WORLD
            world_size = dist.get_world_size(process_group)
        return _SyncBNFunction.apply(input, z, self.weight, self.bias, running_mean, running_var, self.eps,
                                     float(exponential_average_factor), process_group, world_size,
                                     self.fuse_relu, bn_training, nbt)

    @classmethod
    def convert_sync_batchnorm(cls, module, process_group=None):
        """Replace every `BatchNorm*D` (and stock `torch.nn.SyncBatchNorm`) in `module` by this
        class, sharing parameters and buffers, so state dicts stay interchangeable
        (torch/nn/modules/batchnorm.py:844-905; swap site main_supcon.py:224)."""
        module_output = module
        if isinstance(module, _BatchNorm) and not isinstance(module, cls):
            module_output = cls(module.num_features, module.eps, module.momentum, module.affine,
                                module.track_running_stats, process_group)
            if module.affine:
                with torch.no_grad():
                    module_output.weight = module.weight
                    module_output.bias = module.bias
            module_output.running_mean = module.running_mean
            module_output.running_var = module.running_var
            module_output.num_batches_tracked = module.num_batches_tracked
            module_output.training = module.training
            if hasattr(module, 'qconfig'):
                module_output.qconfig = module.qconfig
        for name, child in module.named_children():
            module_output.add_module(name, cls.convert_sync_batchnorm(child, process_group))
        del module
        return module_output
