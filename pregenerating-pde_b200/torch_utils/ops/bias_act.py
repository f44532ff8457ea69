"""bias_act: fused bias + activation.  API of the reference's torch_utils/ops/bias_act.py
(`bias_act(x, b, dim, act, alpha, gain, clamp, impl)`, `activation_funcs`), first and second order
gradients included; kernels: csrc/bias_act.cu."""
import math
import torch

from .. import custom_ops


class _Spec(dict):
    __getattr__ = dict.__getitem__


def _mk(func, def_alpha, def_gain, cuda_idx, ref, has_2nd_grad):
    return _Spec(func=func, def_alpha=def_alpha, def_gain=def_gain, cuda_idx=cuda_idx, ref=ref, has_2nd_grad=has_2nd_grad)


_S2 = math.sqrt(2.0)
# name -> spec; same keys, defaults and cuda_idx as bias_act.py:13-23
activation_funcs = {
    'linear':   _mk(lambda x, **_: x,                                        0,   1,   1, '',  False),
    'relu':     _mk(lambda x, **_: torch.relu(x),                            0,   _S2, 2, 'y', False),
    'lrelu':    _mk(lambda x, alpha, **_: torch.nn.functional.leaky_relu(x, alpha), 0.2, _S2, 3, 'y', False),
    'tanh':     _mk(lambda x, **_: torch.tanh(x),                            0,   1,   4, 'y', True),
    'sigmoid':  _mk(lambda x, **_: torch.sigmoid(x),                         0,   1,   5, 'y', True),
    'elu':      _mk(lambda x, **_: torch.nn.functional.elu(x),               0,   1,   6, 'y', True),
    'selu':     _mk(lambda x, **_: torch.nn.functional.selu(x),              0,   1,   7, 'y', True),
    'softplus': _mk(lambda x, **_: torch.nn.functional.softplus(x),          0,   1,   8, 'y', True),
    'swish':    _mk(lambda x, **_: torch.sigmoid(x) * x,                     0,   _S2, 9, 'x', True),
}

_plugin = None
_EMPTY = torch.empty([0])


def _init():
    global _plugin
    if _plugin is None:
        _plugin = custom_ops.get_plugin('bias_act_plugin')
    return True


def _resolve(act, alpha, gain, clamp):
    assert clamp is None or clamp >= 0
    spec = activation_funcs[act]
    return (spec, float(spec.def_alpha if alpha is None else alpha), float(spec.def_gain if gain is None else gain),
            float(-1 if clamp is None else clamp))


@custom_ops.profiled('cno::bias_act')
def bias_act(x, b=None, dim=1, act='linear', alpha=None, gain=None, clamp=None, impl='cuda'):
    """y = clamp(act(x + b) * gain).  Arguments as in the reference (bias_act.py:44-78)."""
    assert isinstance(x, torch.Tensor)
    assert impl in ['ref', 'cuda']
    if impl == 'cuda' and x.device.type == 'cuda' and _init():
        spec, alpha, gain, clamp = _resolve(act, alpha, gain, clamp)
        return _BiasAct.apply(x, b, dim, act, alpha, gain, clamp)
    return _bias_act_ref(x=x, b=b, dim=dim, act=act, alpha=alpha, gain=gain, clamp=clamp)


def _bias_act_ref(x, b=None, dim=1, act='linear', alpha=None, gain=None, clamp=None):
    """Definition in plain torch ops (reference: bias_act.py:83-112)."""
    assert isinstance(x, torch.Tensor)
    spec, alpha, gain, clamp = _resolve(act, alpha, gain, clamp)
    if b is not None:
        assert isinstance(b, torch.Tensor) and b.ndim == 1
        assert 0 <= dim < x.ndim and b.shape[0] == x.shape[dim]
        view = [1] * x.ndim
        view[dim] = -1
        x = x + b.reshape(view)
    x = spec.func(x, alpha=alpha)
    if gain != 1:
        x = x * gain
    if clamp >= 0:
        x = x.clamp(-clamp, clamp)
    return x


def _fmt(t):
    return torch.channels_last if t.ndim > 2 and t.stride(1) == 1 else torch.contiguous_format


class _BiasAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, b, dim, act, alpha, gain, clamp):
        spec = activation_funcs[act]
        ctx.cfg = (dim, act, alpha, gain, clamp)
        ctx.fmt = _fmt(x)
        x = x.contiguous(memory_format=ctx.fmt)
        bb = b.contiguous() if b is not None else _EMPTY
        y = x
        if act != 'linear' or gain != 1 or clamp >= 0 or b is not None:
            y = _plugin.bias_act(x, bb, _EMPTY, _EMPTY, _EMPTY, 0, dim, spec.cuda_idx, alpha, gain, clamp)
        keep_x = 'x' in spec.ref or spec.has_2nd_grad
        # 'linear' keeps nothing in the reference (ref=''), so its plugin gradient ignores clamp; the impl='ref'
        # definition (the parity target) zeroes the gradient where y was clamped, so y is kept in that case
        keep_y = 'y' in spec.ref or (clamp >= 0 and spec.ref == '')
        ctx.save_for_backward(x if keep_x else _EMPTY, bb if keep_x else _EMPTY, y if keep_y else _EMPTY)
        return y

    @staticmethod
    def backward(ctx, dy):
        dim, act, alpha, gain, clamp = ctx.cfg
        x, b, y = ctx.saved_tensors
        dy = dy.contiguous(memory_format=ctx.fmt)
        dx = db = None
        if ctx.needs_input_grad[0] or ctx.needs_input_grad[1]:
            dx = dy
            if act != 'linear' or gain != 1 or clamp >= 0:
                dx = _BiasActGrad.apply(dy, x, b, y, ctx.cfg)
        if ctx.needs_input_grad[1]:
            db = dx.sum([i for i in range(dx.ndim) if i != dim])
        return dx, db, None, None, None, None, None


class _BiasActGrad(torch.autograd.Function):
    @staticmethod
    def forward(ctx, dy, x, b, y, cfg):
        dim, act, alpha, gain, clamp = cfg
        spec = activation_funcs[act]
        ctx.cfg, ctx.fmt = cfg, _fmt(dy)
        dx = _plugin.bias_act(dy, b, x, y, _EMPTY, 1, dim, spec.cuda_idx, alpha, gain, clamp)
        ctx.save_for_backward(dy if spec.has_2nd_grad else _EMPTY, x, b, y)
        return dx

    @staticmethod
    def backward(ctx, d_dx):
        dim, act, alpha, gain, clamp = ctx.cfg
        spec = activation_funcs[act]
        d_dx = d_dx.contiguous(memory_format=ctx.fmt)
        dy, x, b, y = ctx.saved_tensors
        d_dy = d_x = d_b = None
        if ctx.needs_input_grad[0]:
            d_dy = _BiasActGrad.apply(d_dx, x, b, y, ctx.cfg)
        if spec.has_2nd_grad and (ctx.needs_input_grad[1] or ctx.needs_input_grad[2]):
            d_x = _plugin.bias_act(d_dx, b, x, y, dy, 2, dim, spec.cuda_idx, alpha, gain, clamp)
        if spec.has_2nd_grad and ctx.needs_input_grad[2]:
            d_b = d_x.sum([i for i in range(d_x.ndim) if i != dim])
        return d_dy, d_x, d_b, None, None
