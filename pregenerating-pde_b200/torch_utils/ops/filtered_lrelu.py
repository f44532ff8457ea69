"""filtered_lrelu: bias -> FIR upsample -> gain * leaky ReLU (clamp) -> FIR downsample, one fused
sm_100a kernel (csrc/flrelu_fused*.cu).  API of the reference's torch_utils/ops/filtered_lrelu.py:
same signature, `impl=` switch, sign-mask backward (backward is the same op with up<->down, so
gradients of any order work) and `db = dx.sum([0, 2, 3])`."""
import warnings
import numpy as np
import torch

from .. import custom_ops
from . import upfirdn2d
from . import bias_act

_plugin = None


def _init():
    global _plugin
    if _plugin is None:
        _plugin = custom_ops.get_plugin('filtered_lrelu_plugin')
    return True


def _get_filter_size(f):
    if f is None:
        return 1, 1
    assert isinstance(f, torch.Tensor) and 1 <= f.ndim <= 2
    return f.shape[-1], f.shape[0]   # width, height


def _parse_padding(padding):
    if isinstance(padding, (int, np.integer)):
        padding = [padding, padding]
    assert isinstance(padding, (list, tuple)) and all(isinstance(v, (int, np.integer)) for v in padding)
    padding = [int(v) for v in padding]
    if len(padding) == 2:
        padding = [padding[0], padding[0], padding[1], padding[1]]
    px0, px1, py0, py1 = padding
    return px0, px1, py0, py1


def _check_scalars(up, down, gain, slope, clamp):
    assert isinstance(up, int) and up >= 1
    assert isinstance(down, int) and down >= 1
    assert gain == float(gain) and gain > 0
    assert slope == float(slope) and slope >= 0
    assert clamp is None or (clamp == float(clamp) and clamp >= 0)


@custom_ops.profiled('cno::filtered_lrelu')
def filtered_lrelu(x, fu=None, fd=None, b=None, up=1, down=1, padding=0, gain=np.sqrt(2), slope=0.2, clamp=None,
                   flip_filter=False, impl='cuda'):
    """Arguments and semantics as in the reference (filtered_lrelu.py:48-108):
    x [N,C,H,W]; fu / fd fp32 filters ([taps] separable, [h,w] full, None identity); b [C] of x.dtype;
    padding int | [x,y] | [x0,x1,y0,y1] w.r.t. the upsampled image; returns [N,C,OH,OW]."""
    assert isinstance(x, torch.Tensor)
    assert impl in ['ref', 'cuda']
    if impl == 'cuda' and x.device.type == 'cuda' and _init():
        _check_scalars(up, down, gain, slope, clamp)
        cfg = (up, down) + _parse_padding(padding) + (float(gain), float(slope),
                                                       float('inf') if clamp is None else float(clamp), bool(flip_filter))
        return _FilteredLRelu.apply(x, fu, fd, b, None, 0, 0, cfg, False)
    return _filtered_lrelu_ref(x, fu=fu, fd=fd, b=b, up=up, down=down, padding=padding, gain=gain, slope=slope,
                               clamp=clamp, flip_filter=flip_filter)


def _filtered_lrelu_ref(x, fu=None, fd=None, b=None, up=1, down=1, padding=0, gain=np.sqrt(2), slope=0.2, clamp=None,
                        flip_filter=False):
    """Definition through upfirdn2d / bias_act (reference: filtered_lrelu.py:113-145)."""
    assert isinstance(x, torch.Tensor) and x.ndim == 4
    fu_w, fu_h = _get_filter_size(fu)
    fd_w, fd_h = _get_filter_size(fd)
    if b is not None:
        assert isinstance(b, torch.Tensor) and b.dtype == x.dtype and list(b.shape) == [x.shape[1]]
    _check_scalars(up, down, gain, slope, clamp)
    px0, px1, py0, py1 = _parse_padding(padding)
    n, c, in_h, in_w = x.shape
    out_w = (in_w * up + (px0 + px1) - (fu_w - 1) - (fd_w - 1) + (down - 1)) // down
    out_h = (in_h * up + (py0 + py1) - (fu_h - 1) - (fd_h - 1) + (down - 1)) // down
    in_dtype = x.dtype
    x = bias_act.bias_act(x=x, b=b)
    x = upfirdn2d.upfirdn2d(x=x, f=fu, up=up, padding=[px0, px1, py0, py1], gain=up ** 2, flip_filter=flip_filter)
    x = bias_act.bias_act(x=x, act='lrelu', alpha=slope, gain=gain, clamp=clamp)
    x = upfirdn2d.upfirdn2d(x=x, f=fd, down=down, flip_filter=flip_filter)
    assert list(x.shape) == [n, c, out_h, out_w] and x.dtype == in_dtype
    return x


class _FilteredLRelu(torch.autograd.Function):
    """forward(x, fu, fd, b, si, sx, sy, cfg, signs_only); cfg = (up, down, px0, px1, py0, py1, gain, slope, clamp, flip);
    signs_only: the given mask `si` is known to hold sign codes only (written with clamp = none)."""

    @staticmethod
    def _run(x, fu, fd, b, si, sx, sy, cfg, write_signs, want_db=False, signs_only=False):
        """one fused launch (or the generic composition); returns (y, so, fu, fd, db)"""
        up, down, px0, px1, py0, py1, gain, slope, clamp, flip = cfg
        assert isinstance(x, torch.Tensor) and x.ndim == 4
        if fu is None:
            fu = torch.ones([1, 1], dtype=torch.float32, device=x.device)
        if fd is None:
            fd = torch.ones([1, 1], dtype=torch.float32, device=x.device)
        assert 1 <= fu.ndim <= 2 and 1 <= fd.ndim <= 2
        if up == 1 and fu.ndim == 1 and fu.shape[0] == 1:
            fu = fu.square()[None]
        if down == 1 and fd.ndim == 1 and fd.shape[0] == 1:
            fd = fd.square()[None]
        if si is None:
            si = torch.empty([0])
        no_bias = b is None
        if no_bias:
            b = torch.empty([0], dtype=x.dtype, device=x.device)      # plugin extension: empty = no bias (no zero-fill launch per backward)
        strides = [x.stride(i) for i in range(x.ndim) if x.size(i) > 1]
        if any(a < c for a, c in zip(strides[:-1], strides[1:])):
            warnings.warn('low-performance memory layout detected in filtered_lrelu input', RuntimeWarning)
        # fused kernel (stream-safe: no global filter staging, unlike filtered_lrelu.py:207-208)
        rc = -1
        db = torch.zeros([x.shape[1]], dtype=torch.float32, device=x.device) if want_db else None
        if x.dtype in (torch.float16, torch.bfloat16, torch.float32):
            y, so, rc = _plugin.filtered_lrelu(x, fu, fd, b, si, up, down, px0, px1, py0, py1, sx, sy, gain, slope,
                                               clamp, flip, write_signs, db=db, signs_only=signs_only)
        if rc < 0:
            # generic composition on our own kernels; keeps only the bit-packed mask (filtered_lrelu.py:215-221)
            y = x.clone() if no_bias else x.add(b.unsqueeze(-1).unsqueeze(-1))
            y = upfirdn2d.upfirdn2d(x=y, f=fu, up=up, padding=[px0, px1, py0, py1], gain=up ** 2, flip_filter=flip)
            so = _plugin.filtered_lrelu_act_(y, si, sx, sy, gain, slope, clamp, write_signs)
            y = upfirdn2d.upfirdn2d(x=y, f=fd, down=down, flip_filter=flip)
            if want_db:
                db = y.sum([0, 2, 3])
        if want_db:
            db = db.to(x.dtype)
        return y, (si if si.numel() else so), fu, fd, db

    @staticmethod
    def forward(ctx, x, fu, fd, b, si, sx, sy, cfg, signs_only):
        # the sign mask is produced only when a gradient can be asked for (filtered_lrelu.py:198)
        given = si is not None and si.numel() > 0
        write_signs = (not given) and (x.requires_grad or (b is not None and b.requires_grad))
        y, signs, fu, fd, _ = _FilteredLRelu._run(x, fu, fd, b, si, sx, sy, cfg, write_signs, signs_only=given and signs_only)
        ctx.save_for_backward(fu, fd, signs)
        ctx.x_shape, ctx.y_shape, ctx.s_ofs, ctx.cfg = x.shape, y.shape, (sx, sy), cfg
        # a mask written here holds "clamped" codes only if this call clamps; a given mask keeps what its writer said
        ctx.signs_only = bool(signs_only) if given else (cfg[8] == float('inf'))
        return y

    @staticmethod
    def backward(ctx, dy):
        fu, fd, si = ctx.saved_tensors
        up, down, px0, px1, py0, py1, gain, slope, clamp, flip = ctx.cfg
        _, _, xh, xw = ctx.x_shape
        _, _, yh, yw = ctx.y_shape
        sx, sy = ctx.s_ofs
        assert not any(ctx.needs_input_grad[i] for i in (1, 2, 4, 5, 6))
        dx = db = None
        if ctx.needs_input_grad[0] or ctx.needs_input_grad[3]:
            # swapped-op recipe of filtered_lrelu.py:244-256 (verified in SURVEY Appendix C)
            pp = ((fu.shape[-1] - 1) + (fd.shape[-1] - 1) - px0, xw * up - yw * down + px0 - (up - 1),
                  (fu.shape[0] - 1) + (fd.shape[0] - 1) - py0, xh * up - yh * down + py0 - (up - 1))
            gg = gain * (up ** 2) / (down ** 2)
            cfg = (down, up) + pp + (gg, slope, float('inf'), not flip)
            bsx, bsy = sx - (fu.shape[-1] - 1) + px0, sy - (fu.shape[0] - 1) + py0
            if ctx.needs_input_grad[3] and not (torch.is_grad_enabled() and dy.requires_grad):
                # first-order backward: db comes out of the same kernel (no second pass over dx)
                dx, _, _, _, db = _FilteredLRelu._run(dy, fd, fu, None, si, bsx, bsy, cfg, False, want_db=True, signs_only=ctx.signs_only)
                return dx, None, None, db, None, None, None, None, None
            dx = _FilteredLRelu.apply(dy, fd, fu, None, si, bsx, bsy, cfg, ctx.signs_only)
        if ctx.needs_input_grad[3]:
            db = dx.sum([0, 2, 3])
        return dx, None, None, db, None, None, None, None, None
