"""CIN: the time-conditioned instance norm of the CNO (`FILM`, CNO_timeModule_CIN.py:27-70).

`cin(x, gamma, beta, scale, shift)` is InstanceNorm2d(affine) followed by the per-(n,c) FiLM affine in
one kernel pass (csrc/cin.cu), with a fused backward.  `film_forward` is a drop-in for `FILM.forward`
that keeps the module's parameters / state_dict untouched; the four tiny Linear layers stay in torch
(autograd carries dscale / dshift into them)."""
import ctypes
import torch

from . import _lib
from .torch_utils.custom_ops import _DT, _ptr, _stream


class _Cin(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, scale, shift, eps):
        lib = _lib.load()
        n, c, h, w = x.shape
        x = x.contiguous()
        g32, b32 = gamma.detach().float().contiguous(), beta.detach().float().contiguous()
        s32, q32 = scale.detach().float().contiguous(), shift.detach().float().contiguous()
        y = torch.empty_like(x)
        mean = torch.empty([n, c], dtype=torch.float32, device=x.device)
        rstd = torch.empty_like(mean)
        with torch.cuda.device(x.device):
            rc = lib.cno_cin_fwd(_ptr(x), _ptr(y), _ptr(g32), _ptr(b32), _ptr(s32), _ptr(q32), _ptr(mean), _ptr(rstd),
                                 _DT[x.dtype], n, c, h * w, float(eps), _stream(x))
        _lib.check(rc, 'cin_fwd')
        ctx.save_for_backward(x, g32, b32, s32, mean, rstd)
        ctx.dtypes = (gamma.dtype, beta.dtype, scale.dtype, shift.dtype)
        return y

    @staticmethod
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
        return (dx if ctx.needs_input_grad[0] else None), dg, db, ds, dq, None


def cin(x, gamma, beta, scale, shift, eps=1e-5):
    """x [N,C,H,W] (fp32 / bf16 / fp16, CUDA); gamma, beta [C]; scale, shift [N,C]."""
    assert x.is_cuda and x.ndim == 4
    assert gamma.shape == (x.shape[1],) and beta.shape == (x.shape[1],)
    assert scale.shape == x.shape[:2] and shift.shape == x.shape[:2]
    return _Cin.apply(x, gamma, beta, scale, shift, eps)


def film_forward(self, x, timestep):
    """Replacement for FILM.forward (CNO_timeModule_CIN.py:60-70).  Uses the fused kernel when the module
    is the instance-norm flavour on a CUDA tensor; every other case runs the module's own torch ops."""
    norm = self.norm
    t = timestep.reshape(-1, 1).type_as(x)
    scale = self.lat2scale(self.inp2lat_sacale(t))
    shift = self.lat2bias(self.inp2lat_bias(t))
    fused = (x.is_cuda and x.ndim == 4 and isinstance(norm, torch.nn.InstanceNorm2d) and norm.affine
             and not norm.track_running_stats and x.dtype in _DT and x.dtype != torch.float64)
    if fused:
        return cin(x, norm.weight, norm.bias, scale, shift, norm.eps)
    x = norm(x)
    return x * scale.unsqueeze(2).unsqueeze(3) + shift.unsqueeze(2).unsqueeze(3)
