"""Explicit NCHW <-> NHWC conversion around the cuDNN convolutions (csrc/layout.cu).  cuDNN's tensor-op kernels are NHWC inside; when
they are handed NCHW tensors every pass converts its own operands: x for the forward and again for the weight gradient, dy for the
data gradient and again for the weight gradient, y and dx back - six activation-sized transposes per convolution and training step.
`conv2d_nhwc` converts x once (and keeps the NHWC copy for the backward), dy once, and y / dx back: four, on our own tiled kernel."""
import torch

from . import _lib
from .torch_utils.custom_ops import _ptr, _stream

_DT = {torch.float32: _lib.F32, torch.bfloat16: _lib.BF16, torch.float16: _lib.F16}


def to_nhwc(x):
    """dense NCHW tensor -> the same values with channels_last strides"""
    assert x.ndim == 4 and x.is_cuda and x.dtype in _DT
    x = x.contiguous()
    n, c, h, w = x.shape
    y = torch.empty_like(x, memory_format=torch.channels_last)
    if x.numel():
        _lib.check(_lib.load().cno_nchw_to_nhwc(_ptr(x), _ptr(y), _DT[x.dtype], n, c, h * w, _stream(x)), 'cno_nchw_to_nhwc')
    return y


def to_nchw(x):
    """channels_last tensor -> dense NCHW"""
    assert x.ndim == 4 and x.is_cuda and x.dtype in _DT
    if not x.is_contiguous(memory_format=torch.channels_last):
        return x.contiguous()
    n, c, h, w = x.shape
    y = torch.empty_like(x, memory_format=torch.contiguous_format)
    if x.numel():
        _lib.check(_lib.load().cno_nhwc_to_nchw(_ptr(x), _ptr(y), _DT[x.dtype], n, c, h * w, _stream(x)), 'cno_nhwc_to_nchw')
    return y


class _Conv2dNHWC(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, stride, padding, dilation, groups):
        x_cl = to_nhwc(x)
        w_cl = weight.contiguous(memory_format=torch.channels_last)          # once per step; forward, data and weight gradient all use it
        y_cl = torch.nn.functional.conv2d(x_cl, w_cl, bias, stride, padding, dilation, groups)
        ctx.save_for_backward(x_cl, w_cl)
        ctx.cfg = (stride, padding, dilation, groups, None if bias is None else list(bias.shape))
        return to_nchw(y_cl)

    @staticmethod
    def backward(ctx, dy):
        x_cl, w_cl = ctx.saved_tensors
        stride, padding, dilation, groups, bias_sizes = ctx.cfg
        dy_cl = to_nhwc(dy)
        mask = [ctx.needs_input_grad[0], ctx.needs_input_grad[1], bias_sizes is not None and ctx.needs_input_grad[2]]
        dx_cl, dw, db = torch.ops.aten.convolution_backward(dy_cl, x_cl, w_cl, bias_sizes, list(stride), list(padding), list(dilation), False, [0, 0],
                                                            groups, mask)
        dx = to_nchw(dx_cl) if dx_cl is not None else None
        if dw is not None and not dw.is_contiguous():
            dw = dw.contiguous()
        return dx, dw, db, None, None, None, None


def conv2d_nhwc(x, weight, stride=(1, 1), padding=(0, 0), dilation=(1, 1), groups=1, bias=None):
    """conv2d(x, weight, bias) on NCHW tensors, with the layout conversions cuDNN needs done once each (see module docstring).
    Autocast: runs in the autocast dtype like torch.nn.functional.conv2d would."""
    if torch.is_autocast_enabled('cuda'):
        dt = torch.get_autocast_dtype('cuda')
        x, weight = x.to(dt), weight.to(dt)
        bias = None if bias is None else bias.to(dt)
    return _Conv2dNHWC.apply(x, weight, bias, tuple(stride), tuple(padding), tuple(dilation), groups)
