"""`ProjectionTail` -- the tail of the reference's MLP head as one kernel.

Replaces `F.normalize(self.head[2](h), dim=1)` of /root/reference/networks/resnet_big.py:168-172,177-181 and, with
`n_views=2`, the view stacking of /root/reference/main_supcon.py:285-286 (the second `F.normalize` of :283 is the
identity on unit rows): the output is the `[bsz, n_views, feat_dim]` tensor the loss takes, bf16 rows written by the
GEMM's epilogue -- `SupConLoss` then uses it in place, no prep kernel in between.

    head = model.head                                     # Sequential(Linear, ReLU, Linear)
    tail = ProjectionTail.from_linear(head[2], n_views=2)
    z = tail(head[1](head[0](feat)))                      # feat: encoder output of cat([view0, view1]), [2 * bsz, dim_in]
    loss = criterion(z)                                   # z: [bsz, 2, feat_dim], unit rows

The forward product runs on the tensor cores in bf16 with fp32 accumulation (what autocast runs for a Linear); the
backward is the normalisation's backward (one small kernel) followed by the two GEMMs of a Linear's backward, which are
plain library GEMMs (torch.matmul).  Parameters are named like `nn.Linear`'s (`weight`, `bias`): state dicts are
interchangeable with `head[2]`.
"""
import math

import torch
import torch.nn as nn

from . import _lib
from .losses import _on_device, _stream


class _ProjectionTailFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, h, weight, bias, n_views, normalize, eps, out_dtype):
        lib = _lib.load()
        dev = h.device
        M, K = h.shape
        N = weight.shape[0]
        hb = h.detach()
        if hb.dtype != torch.bfloat16 or hb.stride(1) != 1 or hb.stride(0) % 8 != 0 or hb.data_ptr() % 16 != 0:
            hb = hb.to(torch.bfloat16).contiguous()
        wb = weight.detach().to(torch.bfloat16).contiguous()
        bf = bias.detach().float().contiguous() if bias is not None else None
        out = torch.empty((M, N), dtype=out_dtype, device=dev)
        inv_norm = torch.empty(M, dtype=torch.float32, device=dev)
        with _on_device(dev):
            _lib.check(lib.scb_head_forward(
                hb.data_ptr(), hb.stride(0), wb.data_ptr(), wb.stride(0), bf.data_ptr() if bf is not None else None,
                M, K, N, n_views, out.data_ptr(), _lib.dtype_code(out_dtype), out.stride(0), inv_norm.data_ptr(),
                int(normalize), float(eps), _stream(dev)), 'scb_head_forward')
        ctx.save_for_backward(hb, wb, out, inv_norm)
        ctx.meta = (n_views, normalize, h.dtype, weight.dtype, bias is not None, bias.dtype if bias is not None else None)
        return out.view(M // n_views, n_views, N) if n_views > 1 else out

    @staticmethod
    def backward(ctx, g):
        hb, wb, out, inv_norm = ctx.saved_tensors
        n_views, normalize, h_dtype, w_dtype, has_bias, b_dtype = ctx.meta
        lib = _lib.load()
        dev = g.device
        M, K = hb.shape
        N = wb.shape[0]
        g2 = g.reshape(M, N)
        if g2.stride(1) != 1:
            g2 = g2.contiguous()
        # gradient with respect to the un-normalised product, back in the input's row order
        d_raw = torch.empty((M, N), dtype=torch.bfloat16 if h_dtype == torch.bfloat16 else torch.float32, device=dev)
        if normalize:
            with _on_device(dev):
                _lib.check(lib.scb_head_normalize_backward(
                    g2.data_ptr(), _lib.dtype_code(g2.dtype), g2.stride(0), out.data_ptr(), _lib.dtype_code(out.dtype),
                    out.stride(0), inv_norm.data_ptr(), M, N, d_raw.data_ptr(), _lib.dtype_code(d_raw.dtype),
                    d_raw.stride(0), n_views, _stream(dev)), 'scb_head_normalize_backward')
        else:
            src = g2.view(M // n_views, n_views, N).transpose(0, 1).reshape(M, N) if n_views > 1 else g2
            d_raw.copy_(src)
        # the Linear's own backward: two library GEMMs and a column sum
        grad_h = grad_w = grad_b = None
        if ctx.needs_input_grad[0]:
            grad_h = (d_raw @ wb.to(d_raw.dtype)).to(h_dtype)
        if ctx.needs_input_grad[1]:
            grad_w = (d_raw.t() @ hb.to(d_raw.dtype)).to(w_dtype)
        if has_bias and ctx.needs_input_grad[2]:
            grad_b = d_raw.float().sum(0).to(b_dtype)
        return grad_h, grad_w, grad_b, None, None, None, None


class ProjectionTail(nn.Module):
    """`normalize(Linear(h))`, optionally stacking the views for the loss.  See the module docstring."""

    def __init__(self, in_features, out_features, bias=True, n_views=1, normalize=True, eps=1e-12, out_dtype=None,
                 device=None, dtype=None):
        super().__init__()
        if out_features > 128 or out_features % 4 != 0:
            raise ValueError('ProjectionTail: out_features must be a multiple of 4, at most 128 (got %d)' % out_features)
        if in_features % 8 != 0:
            raise ValueError('ProjectionTail: in_features must be a multiple of 8 (got %d)' % in_features)
        self.in_features, self.out_features = in_features, out_features
        self.n_views, self.normalize, self.eps, self.out_dtype = n_views, normalize, eps, out_dtype
        kw = {'device': device, 'dtype': dtype}
        self.weight = nn.Parameter(torch.empty((out_features, in_features), **kw))
        self.bias = nn.Parameter(torch.empty(out_features, **kw)) if bias else None
        # nn.Linear.reset_parameters
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if self.bias is not None:
            bound = 1 / math.sqrt(in_features)
            nn.init.uniform_(self.bias, -bound, bound)

    @classmethod
    def from_linear(cls, linear, n_views=1, normalize=True, out_dtype=None):
        """Shares the parameters of an existing `nn.Linear` (e.g. `model.head[2]`)."""
        m = cls(linear.in_features, linear.out_features, bias=linear.bias is not None, n_views=n_views,
                normalize=normalize, out_dtype=out_dtype, device=linear.weight.device, dtype=linear.weight.dtype)
        m.weight = linear.weight
        m.bias = linear.bias
        return m

    def forward(self, h):
        if h.dim() != 2 or h.shape[1] != self.in_features:
            raise ValueError('ProjectionTail: expected input [rows, %d], got %s' % (self.in_features, tuple(h.shape)))
        if not h.is_cuda:
            raise RuntimeError('simclr_pytorch_distributed_b200.ProjectionTail runs on CUDA tensors only '
                               '(there is no CPU fallback); got a %s tensor' % h.device.type)
        if h.shape[0] % self.n_views != 0:
            raise ValueError('ProjectionTail: %d rows are not a multiple of n_views=%d' % (h.shape[0], self.n_views))
        out_dtype = self.out_dtype or (torch.bfloat16 if h.dtype == torch.bfloat16 else
                                       (torch.float16 if h.dtype == torch.float16 else torch.float32))
        return _ProjectionTailFunction.apply(h, self.weight, self.bias, self.n_views, self.normalize, self.eps, out_dtype)

    def extra_repr(self):
        return 'in_features=%d, out_features=%d, bias=%s, n_views=%d, normalize=%s' % (
            self.in_features, self.out_features, self.bias is not None, self.n_views, self.normalize)
