"""CIN: the time-conditioned instance norm of the CNO (`FILM`, CNO_timeModule_CIN.py:27-70).

`cin(x, gamma, beta, scale, shift)` is InstanceNorm2d(affine) followed by the per-(n,c) FiLM affine in
one kernel pass (csrc/cin.cu), with a fused backward.  `film_forward` is a drop-in for `FILM.forward`
that keeps the module's parameters / state_dict untouched; the four tiny Linear layers stay in torch
(autograd carries dscale / dshift into them)."""
import contextlib
import ctypes
import os
import torch

from . import _lib
from .torch_utils.custom_ops import _DT, _ptr, _stream, profiled as _profiled


FUSE_LRELU = os.environ.get('CNO_FUSE_LRELU', '1') != '0'         # False: film_lrelu runs FILM and the activation as two ops (the normalised tensor goes through HBM)
FUSE_LINEARS = True       # False: keep the four Linear layers in torch (autograd supports higher-order gradients there)


class _Cin(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, scale, shift, eps, res=None):
        lib = _lib.load()
        n, c, h, w = x.shape
        x = x.contiguous()
        res = None if res is None else res.to(x.dtype).contiguous()
        g32, b32 = gamma.detach().float().contiguous(), beta.detach().float().contiguous()
        s32, q32 = scale.detach().float().contiguous(), shift.detach().float().contiguous()
        y = torch.empty_like(x)
        mean = torch.empty([n, c], dtype=torch.float32, device=x.device)
        rstd = torch.empty_like(mean)
        with torch.cuda.device(x.device):
            rc = lib.cno_cin_fwd_res(_ptr(x), _ptr(res), _ptr(y), _ptr(g32), _ptr(b32), _ptr(s32), _ptr(q32), _ptr(mean), _ptr(rstd),
                                 _DT[x.dtype], n, c, h * w, float(eps), _stream(x))
        _lib.check(rc, 'cin_fwd')
        ctx.save_for_backward(x, g32, b32, s32, mean, rstd)
        ctx.dtypes = (gamma.dtype, beta.dtype, scale.dtype, shift.dtype)
        return y

    @staticmethod
    @torch.autograd.function.once_differentiable      # dx comes from an opaque kernel: ask for create_graph and it raises instead of dropping terms
    def backward(ctx, dy):
        lib = _lib.load()
        x, g32, b32, s32, mean, rstd = ctx.saved_tensors
        n, c, h, w = x.shape
        dy = dy.contiguous()
        dx = torch.empty_like(x)
        s1 = torch.empty([n, c], dtype=torch.float32, device=x.device)
        s2 = torch.empty_like(s1)
        with torch.cuda.device(x.device):
            rc = lib.cno_cin_bwd(_ptr(x), _ptr(dy), _ptr(dx), _ptr(g32), _ptr(s32), _ptr(mean), _ptr(rstd), _ptr(s1), _ptr(s2),
                                 _DT[x.dtype], n, c, h * w, _stream(x))
        _lib.check(rc, 'cin_bwd')
        dg = db = ds = dq = None
        gd, bd, sd, qd = ctx.dtypes
        if ctx.needs_input_grad[1]:
            dg = (s32 * s2).sum(0).to(gd)
        if ctx.needs_input_grad[2]:
            db = (s32 * s1).sum(0).to(bd)
        if ctx.needs_input_grad[3]:
            ds = (g32 * s2 + b32 * s1).to(sd)
        if ctx.needs_input_grad[4]:
            dq = s1.to(qd)
        return (dx if ctx.needs_input_grad[0] else None), dg, db, ds, dq, None, (dy if ctx.needs_input_grad[6] else None)


def cin(x, gamma, beta, scale, shift, eps=1e-5, residual=None):
    """x [N,C,H,W] (fp32 / bf16 / fp16, CUDA); gamma, beta [C]; scale, shift [N,C]; residual (optional, shape of x): returns
    residual + CIN(x) from the same pass."""
    assert residual is None or residual.shape == x.shape
    assert x.is_cuda and x.ndim == 4
    assert gamma.shape == (x.shape[1],) and beta.shape == (x.shape[1],)
    assert scale.shape == x.shape[:2] and shift.shape == x.shape[:2]
    return _Cin.apply(x, gamma, beta, scale, shift, eps, residual)


class _Film(torch.autograd.Function):
    """FILM.forward as one kernel (scale / shift from the four Linear layers computed in the kernel prologue) and a
    backward of two kernels (cno_cin_bwd + cno_film_bwd_finish) instead of ~30 small torch launches."""

    @staticmethod
    def forward(ctx, x, t, gamma, beta, w1s, b1s, W2s, b2s, w1b, b1b, W2b, b2b, eps, res=None):
        lib = _lib.load()
        n, c, h, w = x.shape
        x = x.contiguous()
        res = None if res is None else res.to(x.dtype).contiguous()
        t32, P, lat = _film_params(t, gamma, beta, w1s, b1s, W2s, b2s, w1b, b1b, W2b, b2b, n, c)
        y = torch.empty_like(x)
        mean = torch.empty([n, c], dtype=torch.float32, device=x.device)
        rstd, scale = torch.empty_like(mean), torch.empty_like(mean)
        with torch.cuda.device(x.device):
            rc = lib.cno_film_fwd_res(_ptr(x), _ptr(res), _ptr(y), _ptr(t32), *[_ptr(v) for v in P], _ptr(scale), _ptr(mean), _ptr(rstd),
                                  _DT[x.dtype], n, c, h * w, lat, float(eps), _stream(x))
        _lib.check(rc, 'film_fwd')
        ctx.save_for_backward(x, t32, scale, mean, rstd, *P)
        ctx.meta = [(v.dtype, v.shape) for v in (gamma, beta, w1s, b1s, W2s, b2s, w1b, b1b, W2b, b2b)]
        return y

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, dy):
        dy = dy.contiguous()
        dx, outs = _film_backward(ctx.saved_tensors, dy, ctx.meta)
        outs = [o if need else None for o, need in zip(outs, ctx.needs_input_grad[2:12])]
        return (dx if ctx.needs_input_grad[0] else None, None, *outs, None, dy if ctx.needs_input_grad[13] else None)


def _film_params(t, gamma, beta, w1s, b1s, W2s, b2s, w1b, b1b, W2b, b2b, n, c):
    f32 = lambda v: v.detach().float().contiguous()
    t32 = f32(t).reshape(-1)
    P = [f32(v) for v in (gamma, beta, w1s.reshape(-1), b1s, W2s, b2s, w1b.reshape(-1), b1b, W2b, b2b)]
    lat = P[2].numel()
    assert t32.numel() == n and P[4].shape == (c, lat) and P[8].shape == (c, lat)
    return t32, P, lat


# Parameter-gradient side stream (opt-in, set by train.Trainer): cno_film_bwd_finish turns the [N, C] sums of a FILM backward into
# the module's ten parameter gradients.  Nothing in backward consumes them, yet at 8.7 us of dependent L2 round trips x 97 modules
# they are 0.85 ms on the critical path of the training step.  With a side stream they run next to the remaining backward (the
# launch occupies 6-36 CTAs); the consumer (bucket gather / optimiser) joins with join_param_grad_stream().  None: same stream.
PARAM_GRAD_STREAM = None
_side_keepalive = []          # buffers read / written on the side stream, kept until the join (no allocator reuse under a running kernel)


def join_param_grad_stream():
    if PARAM_GRAD_STREAM is not None and _side_keepalive:
        torch.cuda.current_stream(PARAM_GRAD_STREAM.device).wait_stream(PARAM_GRAD_STREAM)
        _side_keepalive.clear()


def _film_backward(saved, dy, meta):
    """dx and the ten parameter gradients of one FILM module from (x, t, scale, mean, rstd, weights) and d(FILM output):
    cno_cin_bwd + cno_film_bwd_finish."""
    lib = _lib.load()
    x, t32, scale, mean, rstd, gamma, beta, w1s, b1s, W2s, b2s, w1b, b1b, W2b, b2b = saved
    n, c, h, w = x.shape
    lat = w1s.numel()
    dx = torch.empty_like(x)
    s12 = torch.empty([2, n, c], dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        rc = lib.cno_cin_bwd(_ptr(x), _ptr(dy), _ptr(dx), _ptr(gamma), _ptr(scale), _ptr(mean), _ptr(rstd), _ptr(s12[0]), _ptr(s12[1]),
                             _DT[x.dtype], n, c, h * w, _stream(x))
        _lib.check(rc, 'cin_bwd')
        # one flat buffer for the ten parameter gradients
        sizes = [c, c, lat, lat, c * lat, c, lat, lat, c * lat, c]      # dgamma dbeta dw1s db1s dW2s db2s dw1b db1b dW2b db2b
        flat = torch.empty([sum(sizes)], dtype=torch.float32, device=x.device)
        g = list(flat.split(sizes))
        side = PARAM_GRAD_STREAM if (PARAM_GRAD_STREAM is not None and PARAM_GRAD_STREAM.device == x.device) else None
        if side is not None:
            side.wait_stream(torch.cuda.current_stream(x.device))       # the sums of cno_cin_bwd
            _side_keepalive.append((s12, flat, saved))
        with (torch.cuda.stream(side) if side is not None else contextlib.nullcontext()):
            rc = lib.cno_film_bwd_finish(_ptr(s12[0]), _ptr(s12[1]), _ptr(t32), _ptr(gamma), _ptr(beta), _ptr(scale),
                                         _ptr(w1s), _ptr(b1s), _ptr(W2s), _ptr(w1b), _ptr(b1b), _ptr(W2b),
                                         _ptr(g[0]), _ptr(g[1]), _ptr(g[4]), _ptr(g[5]), _ptr(g[2]), _ptr(g[3]),
                                         _ptr(g[8]), _ptr(g[9]), _ptr(g[6]), _ptr(g[7]), n, c, lat, _stream(x))
        _lib.check(rc, 'film_bwd_finish')
    return dx, [gi.view(shape).to(dt) for gi, (dt, shape) in zip(g, meta)]


class _FilmLRelu(torch.autograd.Function):
    """filtered_lrelu(FILM(x, t)) with the normalised tensor never written: cno_film_stats leaves the plane affine,
    cno_filtered_lrelu_affine applies it while the up-FIR runs.  Backward: the usual SIGN_READ launch on dy gives d(FILM output) and
    the LReLu bias gradient, then the FILM backward as in _Film.  cfg as in _FilteredLRelu (clamp = none)."""

    @staticmethod
    def forward(ctx, x, t, gamma, beta, w1s, b1s, W2s, b2s, w1b, b1b, W2b, b2b, eps, fu, fd, b, cfg):
        from .torch_utils.ops import filtered_lrelu as flr
        lib = _lib.load()
        flr._init()
        n, c, h, w = x.shape
        x = x.contiguous()
        t32, P, lat = _film_params(t, gamma, beta, w1s, b1s, W2s, b2s, w1b, b1b, W2b, b2b, n, c)
        mean = torch.empty([n, c], dtype=torch.float32, device=x.device)
        rstd, scale, aff_a, aff_d = (torch.empty_like(mean) for _ in range(4))
        with torch.cuda.device(x.device):
            rc = lib.cno_film_stats(_ptr(x), _ptr(t32), *[_ptr(v) for v in P], _ptr(scale), _ptr(mean), _ptr(rstd), _ptr(aff_a), _ptr(aff_d),
                                    _DT[x.dtype], n, c, h * w, lat, float(eps), _stream(x))
        _lib.check(rc, 'film_stats')
        up, down, px0, px1, py0, py1, gain, slope, clamp, flip = cfg
        write_signs = any(ctx.needs_input_grad)
        y, so, rc = flr._plugin.filtered_lrelu(x, fu, fd, b.to(x.dtype), torch.empty([0]), up, down, px0, px1, py0, py1, 0, 0, gain, slope,
                                               clamp, flip, write_signs, affine=(aff_a, aff_d))
        if rc < 0:
            raise RuntimeError('film_lrelu: configuration has no fused kernel (check film_lrelu_fusable first)')
        ctx.save_for_backward(x, t32, scale, mean, rstd, *P, fu, fd, so)
        ctx.meta = [(v.dtype, v.shape) for v in (gamma, beta, w1s, b1s, W2s, b2s, w1b, b1b, W2b, b2b)]
        ctx.cfg, ctx.y_shape, ctx.b_dtype = cfg, y.shape, b.dtype
        return y

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, dy):
        from .torch_utils.ops import filtered_lrelu as flr
        *film_saved, fu, fd, si = ctx.saved_tensors
        up, down, px0, px1, py0, py1, gain, slope, clamp, flip = ctx.cfg
        _, _, xh, xw = film_saved[0].shape
        _, _, yh, yw = ctx.y_shape
        # swapped-op recipe of filtered_lrelu.py:244-256, as in _FilteredLRelu.backward
        pp = ((fu.shape[-1] - 1) + (fd.shape[-1] - 1) - px0, xw * up - yw * down + px0 - (up - 1),
              (fu.shape[0] - 1) + (fd.shape[0] - 1) - py0, xh * up - yh * down + py0 - (up - 1))
        bcfg = (down, up) + pp + (gain * (up ** 2) / (down ** 2), slope, float('inf'), not flip)
        bsx, bsy = -(fu.shape[-1] - 1) + px0, -(fu.shape[0] - 1) + py0
        dz, _, _, _, db = flr._FilteredLRelu._run(dy.contiguous(), fd, fu, None, si, bsx, bsy, bcfg, False, want_db=True, signs_only=True)
        dx, outs = _film_backward(film_saved, dz, ctx.meta)
        outs = [o if need else None for o, need in zip(outs, ctx.needs_input_grad[2:12])]
        return (dx if ctx.needs_input_grad[0] else None, None, *outs, None, None, None,
                db.to(ctx.b_dtype) if ctx.needs_input_grad[15] else None, None)


def _film_fusable(self, x):
    lin = (self.inp2lat_sacale, self.lat2scale, self.inp2lat_bias, self.lat2bias)
    return (all(isinstance(m, torch.nn.Linear) and m.bias is not None for m in lin)
            and self.inp2lat_sacale.in_features == 1 and self.inp2lat_bias.in_features == 1
            and self.inp2lat_sacale.out_features == self.inp2lat_bias.out_features
            and all(m.weight.dtype == torch.float32 for m in lin))


@_profiled('cno::film')
def film_forward(self, x, timestep, residual=None):
    """Replacement for FILM.forward (CNO_timeModule_CIN.py:60-70).  Uses the fused kernel when the module
    is the instance-norm flavour on a CUDA tensor; every other case runs the module's own torch ops.
    residual (extension): returns residual + FILM(x), added in the kernel's single pass (ResidualBlock epilogue, :503-507)."""
    norm = self.norm
    fused = (x.is_cuda and x.ndim == 4 and isinstance(norm, torch.nn.InstanceNorm2d) and norm.affine
             and not norm.track_running_stats and x.dtype in _DT and x.dtype != torch.float64)
    if fused and FUSE_LINEARS and _film_fusable(self, x) and not timestep.requires_grad:
        return _Film.apply(x, timestep, norm.weight, norm.bias, self.inp2lat_sacale.weight, self.inp2lat_sacale.bias,
                           self.lat2scale.weight, self.lat2scale.bias, self.inp2lat_bias.weight, self.inp2lat_bias.bias,
                           self.lat2bias.weight, self.lat2bias.bias, norm.eps, residual)
    t = timestep.reshape(-1, 1).to(self.inp2lat_sacale.weight.dtype)     # the reference casts to x.dtype (:63), which only works for fp32 x
    scale = self.lat2scale(self.inp2lat_sacale(t))
    shift = self.lat2bias(self.inp2lat_bias(t))
    if fused:
        return cin(x, norm.weight, norm.bias, scale, shift, norm.eps, residual)
    x = norm(x)
    y = x * scale.unsqueeze(2).unsqueeze(3) + shift.unsqueeze(2).unsqueeze(3)
    return y if residual is None else residual + y


SKIP_DEAD_BIAS = os.environ.get('CNO_KEEP_CONV_BIAS', '0') == '0'
CONV_CHANNELS_LAST = os.environ.get('CNO_CONV_CHANNELS_LAST', '1') == '1'


def conv_before_film(conv, film, x):
    """conv(x) for a convolution whose output goes straight into `film` (CNOBlock.forward, CNO_timeModule_CIN.py:329-333;
    ResidualBlock.forward :496-505).  An instance norm subtracts the per-plane mean, so a per-channel constant added in front of
    it vanishes exactly: InstanceNorm(conv(x) + b) == InstanceNorm(conv(x)), and d loss / d b == 0.  The reference still runs
    the bias add (one elementwise launch) and, in backward, the reduction that produces a gradient of pure round-off noise;
    at the TrainCNO_time_L configuration that is 97 x 2 launches and 2.8 ms of the 28 ms training step.  Here the convolution
    runs without the bias and the parameter simply gets no gradient (AdamW skips it; its value stays what the checkpoint
    says).  Only for the per-plane flavour of FILM (nl_dim = [2, 3], no running statistics); CNO_KEEP_CONV_BIAS=1 restores the add."""
    norm = getattr(film, 'norm', None)
    if (SKIP_DEAD_BIAS and conv.bias is not None and isinstance(norm, torch.nn.InstanceNorm2d) and not norm.track_running_stats
            and isinstance(conv, torch.nn.Conv2d) and type(conv).forward is torch.nn.Conv2d.forward and conv.padding_mode == 'zeros'):
        if CONV_CHANNELS_LAST and x.is_cuda and isinstance(conv.padding, tuple) and x.dtype in _DT and x.dtype != torch.float64:
            # cuDNN's tensor-op convolutions are NHWC inside: fed NCHW tensors, every pass converts its operands itself - x for the
            # forward AND again for the weight gradient, dy for the data AND again for the weight gradient, y and dx back - six
            # activation-sized transposes per convolution and training step (7.5 ms of the step, section 6).  With the conversion made
            # explicit (layout.py, csrc/layout.cu) the NHWC x is kept for the backward and dy is converted once: four transposes.
            from . import layout
            return layout.conv2d_nhwc(x, conv.weight, conv.stride, conv.padding, conv.dilation, conv.groups)
        return torch.nn.functional.conv2d(x, conv.weight, None, conv.stride, conv.padding, conv.dilation, conv.groups)
    return conv(x)


def conv_nhwc(conv, x):
    """conv(x) for a plain nn.Conv2d that keeps its bias (LiftProjectBlock's output convolution, CNO_timeModule_CIN.py:163-166): on
    CUDA through layout.conv2d_nhwc, i.e. with the layout conversions cuDNN needs done once each instead of per pass."""
    if (CONV_CHANNELS_LAST and x.is_cuda and isinstance(conv, torch.nn.Conv2d) and type(conv).forward is torch.nn.Conv2d.forward
            and conv.padding_mode == 'zeros' and isinstance(conv.padding, tuple) and x.ndim == 4 and x.dtype in _DT and x.dtype != torch.float64):
        from . import layout
        return layout.conv2d_nhwc(x, conv.weight, conv.stride, conv.padding, conv.dilation, conv.groups, bias=conv.bias)
    return conv(x)


def _film_kernel_ok(self, x, timestep):
    norm = self.norm
    return (x.is_cuda and x.ndim == 4 and isinstance(norm, torch.nn.InstanceNorm2d) and norm.affine
            and not norm.track_running_stats and x.dtype in _DT and x.dtype != torch.float64
            and FUSE_LINEARS and _film_fusable(self, x) and not timestep.requires_grad)


@_profiled('cno::film_lrelu')
def film_lrelu(film, act, x, timestep):
    """act(film(x, timestep)) for a FILM module followed by an LReLu module (CNOBlock.forward, CNO_timeModule_CIN.py:329-335;
    ResidualBlock.forward :499-500).  When both have a fused kernel the normalised tensor is never written: one statistics pass
    over x, then the filtered_lrelu kernel applies the plane affine in its prologue.  Anything else runs the two modules."""
    fu, fd = getattr(act, 'up_filter', None), getattr(act, 'down_filter', None)
    ok = (FUSE_LRELU and x.numel() > 0 and type(film).forward is film_forward and _film_kernel_ok(film, x, timestep) and getattr(act, 'impl', 'cuda') == 'cuda'
          and isinstance(fu, torch.Tensor) and isinstance(fd, torch.Tensor) and fu.ndim == 1 and fd.ndim == 1
          and fu.dtype == torch.float32 and fd.dtype == torch.float32 and type(act).forward is _lrelu_forward_of(act))
    if ok:
        up, down = int(act.up_factor), int(act.down_factor)
        px0, px1, py0, py1 = [int(v) for v in act.padding]
        n, c, xh, xw = x.shape
        fut, fdt = fu.shape[0] - 1, fd.shape[0] - 1
        cw, ch = xw * up + px0 + px1 - fut, xh * up + py0 + py1 - fut
        ok = cw > fdt and ch > fdt
        if ok:
            yw, yh = (cw - fdt + down - 1) // down, (ch - fdt + down - 1) // down
            ok = _lib.load().cno_filtered_lrelu_query(_DT[x.dtype], xh, xw, yh, yw, fu.shape[0], fd.shape[0], up, down, px0, py0) == 0
    if not ok:
        return act(film(x, timestep))
    norm = film.norm
    cfg = (up, down, px0, px1, py0, py1, float(2.0 ** 0.5), 0.2, float('inf'), False)
    return _FilmLRelu.apply(x, timestep, norm.weight, norm.bias, film.inp2lat_sacale.weight, film.inp2lat_sacale.bias,
                            film.lat2scale.weight, film.lat2scale.bias, film.inp2lat_bias.weight, film.inp2lat_bias.bias,
                            film.lat2bias.weight, film.lat2bias.bias, norm.eps, fu, fd, act.bias, cfg)


_LRELU_FORWARDS = {}


def register_lrelu(cls):
    """Declare that cls.forward(x) is exactly filtered_lrelu(x, fu=up_filter, fd=down_filter, b=bias, up=up_factor, down=down_factor,
    padding=padding, gain=sqrt(2), slope=0.2, clamp=None) (training/filtered_networks.py:267-276) so film_lrelu may replace it."""
    _LRELU_FORWARDS[cls] = cls.forward
    return cls


def _lrelu_forward_of(act):
    return _LRELU_FORWARDS.get(type(act))
