"""Explicit NCHW <-> NHWC conversion around the cuDNN convolutions (csrc/layout.cu).  cuDNN's tensor-op kernels are NHWC inside; when
they are handed NCHW tensors every pass converts its own operands: x for the forward and again for the weight gradient, dy for the
data gradient and again for the weight gradient, y and dx back - six activation-sized transposes per convolution and training step.
`conv2d_nhwc` converts x once (and keeps the NHWC copy for the backward), dy once, and y / dx back: four, on our own tiled kernel."""
import contextlib
import os

import torch
from torch.autograd.function import once_differentiable

from . import _lib
from .torch_utils.custom_ops import _ptr, _stream

WGRAD_SIDE_STREAM = os.environ.get('CNO_WGRAD_SIDE_STREAM', '1') == '1'
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
    def forward(ctx, x, weight, bias, stride, padding, dilation, groups, cast):
        x_cl = to_nhwc(x)
        ctx.param_dtypes = (weight.dtype, None if bias is None else bias.dtype)
        ctx.weight_stride, ctx.weight_contiguous = weight.stride(), weight.is_contiguous()
        if cast is not None:                             # autocast: parameters are cast HERE so that their gradients are cast back on the side stream
            bias = None if bias is None else bias.to(cast)
        # once per step, cast and layout in one copy; forward, data gradient and weight gradient all use it
        w_cl = weight.to(dtype=cast if cast is not None else weight.dtype, memory_format=torch.channels_last)
        y_cl = torch.nn.functional.conv2d(x_cl, w_cl, bias, stride, padding, dilation, groups)
        ctx.save_for_backward(x_cl, w_cl)
        ctx.cfg = (stride, padding, dilation, groups, None if bias is None else list(bias.shape))
        return to_nchw(y_cl)

    @staticmethod
    @once_differentiable                                 # first-order only (the training loops need no more); a second backward raises
    def backward(ctx, dy):
        x_cl, w_cl = ctx.saved_tensors
        stride, padding, dilation, groups, bias_sizes = ctx.cfg
        dy_cl = to_nhwc(dy)
        args = (list(stride), list(padding), list(dilation), False, [0, 0], groups)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:                      # the data gradient is what the rest of backward waits for
            dx_cl = torch.ops.aten.convolution_backward(dy_cl, x_cl, w_cl, bias_sizes, *args, [True, False, False])[0]
            dx = to_nchw(dx_cl)
        want_b = bias_sizes is not None and ctx.needs_input_grad[2]
        if ctx.needs_input_grad[1] or want_b:
            # the weight gradient (a quarter of the training step in cuDNN's wgrad kernels) is needed only by the optimiser: it runs on
            # the parameter-gradient side stream (cin.PARAM_GRAD_STREAM, set by the Trainer and joined before gradients are used),
            # under the activation / CIN kernels of the layers below
            from . import cin as _cin
            side = _cin.PARAM_GRAD_STREAM if (_cin.PARAM_GRAD_STREAM is not None and _cin.PARAM_GRAD_STREAM.device == dy.device and WGRAD_SIDE_STREAM) else None
            if side is not None:
                side.wait_stream(torch.cuda.current_stream(dy.device))
                _cin._side_keepalive.append((dy_cl, x_cl, w_cl))
            with (torch.cuda.stream(side) if side is not None else contextlib.nullcontext()):
                _, dw, db = torch.ops.aten.convolution_backward(dy_cl, x_cl, w_cl, bias_sizes, *args, [False, ctx.needs_input_grad[1], want_b])
                if dw is not None:                       # in the layout of the parameter (channels_last weights get their gradient as cuDNN wrote it)
                    if dw.dtype != ctx.param_dtypes[0] or any(sa != sb for n, sa, sb in zip(dw.shape, dw.stride(), ctx.weight_stride) if n > 1):
                        dw = dw.to(dtype=ctx.param_dtypes[0],            # cast and layout in one copy
                                   memory_format=torch.contiguous_format if ctx.weight_contiguous else torch.channels_last)
                if db is not None:
                    db = db.to(ctx.param_dtypes[1])
                # dw / db are NOT put on the keep-alive list: autograd must be able to adopt them as .grad without a copy (a second
                # reference would make AccumulateGrad clone them on the main stream, before the side stream has written them)
        return dx, dw, db, None, None, None, None, None


def conv2d_nhwc(x, weight, stride=(1, 1), padding=(0, 0), dilation=(1, 1), groups=1, bias=None):
    """conv2d(x, weight, bias) on NCHW tensors, with the layout conversions cuDNN needs done once each (see module docstring).
    Autocast: runs in the autocast dtype like torch.nn.functional.conv2d would.  First-order gradients only.  When
    `cin.PARAM_GRAD_STREAM` is set (the Trainer does), weight / bias gradients are produced on that stream: whoever set it calls
    `cin.join_param_grad_stream()` before reading them (Trainer: before the gradient all-reduce and before AdamW)."""
    if torch.is_autocast_enabled('cuda'):
        dt = torch.get_autocast_dtype('cuda')
        return _Conv2dNHWC.apply(x.to(dt), weight, bias, tuple(stride), tuple(padding), tuple(dilation), groups, dt)
    return _Conv2dNHWC.apply(x, weight, bias, tuple(stride), tuple(padding), tuple(dilation), groups, None)
