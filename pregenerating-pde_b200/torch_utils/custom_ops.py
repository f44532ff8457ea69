"""Plugin loader of the B200 build.

The reference's `custom_ops.get_plugin(module_name, sources, ...)` (torch_utils/custom_ops.py:53-149)
JIT-compiles a pybind11 extension per op and returns the module.  Here `get_plugin` returns objects with
the SAME callables (`bias_act`, `upfirdn2d`, `filtered_lrelu`, `filtered_lrelu_act_`, argument for
argument as in bias_act.cpp:24, upfirdn2d.cpp:8, filtered_lrelu.cpp:8/:205) bound to the prebuilt C-ABI
library, so the reference's own op wrappers run on top of it unchanged, and so do ours.
"""
import ctypes
import weakref
import numpy as np
import os
import torch

from .. import _lib

verbosity = 'brief'   # kept for API parity (custom_ops.py:18); nothing is compiled at run time

_DT = {torch.float32: _lib.F32, torch.bfloat16: _lib.BF16, torch.float16: _lib.F16, torch.float64: _lib.F64}


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if (t is not None and t.numel() > 0) else None


def _stream(t):
    return ctypes.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def _dense(t):
    """non-overlapping and dense in some dimension order (bias_act.cpp:41 has_same_layout / is_non_overlapping_and_dense)"""
    if t.is_contiguous() or t.is_contiguous(memory_format=torch.channels_last if t.ndim == 4 else torch.contiguous_format):
        return True
    dims = sorted((d for d in range(t.ndim) if t.size(d) > 1), key=lambda d: t.stride(d))
    expect = 1
    for d in dims:
        if t.stride(d) != expect:
            return False
        expect *= t.size(d)
    return True


def _same_layout(a, b):
    """bias_act.cpp:8-20 has_same_layout: strides are compared only on dimensions of size >= 2"""
    return all(a.size(i) == b.size(i) and (a.size(i) < 2 or a.stride(i) == b.stride(i)) for i in range(a.ndim))


_host_taps = {}     # id(tensor) -> (weakref to the tensor, version, host copy)


def host_taps(f):
    """fp32 taps of a (constant) device filter on the host, so the hot path never synchronises.
    The copy is cached per tensor OBJECT (identity checked through a weak reference whose callback drops the entry)
    together with its version counter: a freed filter whose storage the caching allocator hands to the next filter can
    never alias an entry (an address-keyed cache did), and an in-place update of the filter invalidates the copy."""
    key = id(f)
    hit = _host_taps.get(key)
    if hit is not None and hit[0]() is f and hit[1] == f._version:
        return hit[2]
    taps = np.ascontiguousarray(f.detach().to('cpu', torch.float32).numpy())
    _host_taps[key] = (weakref.ref(f, lambda _r, k=key: _host_taps.pop(k, None)), f._version, taps)
    return taps


# ---- profiler ranges (the reference wraps its ops in torch.autograd.profiler.record_function, torch_utils/misc.py:92-97)
PROFILE_RANGES = bool(os.environ.get('CNO_NVTX'))      # off by default: a range costs host time on every call of the eager path


def profiled(name):
    """decorator: NVTX range + autograd-profiler record around an op entry point when CNO_NVTX=1 (or custom_ops.PROFILE_RANGES = True)"""
    def wrap(fn):
        def inner(*a, **k):
            if not PROFILE_RANGES:
                return fn(*a, **k)
            with torch.autograd.profiler.record_function(name):
                if torch.cuda.is_available():
                    torch.cuda.nvtx.range_push(name)
                try:
                    return fn(*a, **k)
                finally:
                    if torch.cuda.is_available():
                        torch.cuda.nvtx.range_pop()
        inner.__name__, inner.__doc__, inner.__wrapped__ = fn.__name__, fn.__doc__, fn
        return inner
    return wrap


class _BiasActPlugin:
    @staticmethod
    def bias_act(x, b, xref, yref, dy, grad, dim, act, alpha, gain, clamp):
        """bias_act.cpp:24-82.  Empty tensors mean "not given"."""
        lib = _lib.load()
        if not x.is_cuda:
            raise RuntimeError('x must reside on CUDA device')
        if x.dtype not in _DT:
            raise RuntimeError('x must be float16, bfloat16, float32 or float64')
        for name, t in (('b', b), ('xref', xref), ('yref', yref), ('dy', dy)):
            if t.numel() and t.dtype != x.dtype:
                raise RuntimeError(f'{name} must have the same dtype as x')
        for name, t in (('xref', xref), ('yref', yref), ('dy', dy)):
            if t.numel() and (t.shape != x.shape or not _same_layout(t, x)):
                raise RuntimeError(f'{name} must have the same shape and layout as x')
        if b.numel() and (b.ndim != 1 or not (0 <= dim < x.ndim) or b.shape[0] != x.shape[dim]):
            raise RuntimeError('b must be a vector matching dimension `dim` of x')
        if x.numel() > 2 ** 31 - 1:
            raise RuntimeError('x is too large')
        if not _dense(x):
            raise RuntimeError('x must be non-overlapping and dense')
        if grad < 0:
            raise RuntimeError('grad must be non-negative')
        y = torch.empty_like(x)
        if x.numel() == 0:
            return y
        b = b.contiguous()
        with torch.cuda.device(x.device):
            rc = lib.cno_bias_act(_ptr(x), _ptr(b), _ptr(xref), _ptr(yref), _ptr(dy), _ptr(y), _DT[x.dtype], x.numel(),
                                  max(b.numel(), 1), x.stride(dim) if b.numel() else 1, int(grad), int(act),
                                  float(alpha), float(gain), float(clamp), _stream(x))
        _lib.check(rc, 'bias_act')
        return y


class _Upfirdn2dPlugin:
    @staticmethod
    def upfirdn2d(x, f, upx, upy, downx, downy, padx0, padx1, pady0, pady1, flip, gain):
        """upfirdn2d.cpp:8-90 (f is rank 2: [fh, fw])."""
        return _upfirdn2d_call(x, f, False, upx, upy, downx, downy, padx0, padx1, pady0, pady1, flip, gain)


def _upfirdn2d_call(x, f, separable, upx, upy, downx, downy, padx0, padx1, pady0, pady1, flip, gain):
    lib = _lib.load()
    if not x.is_cuda:
        raise RuntimeError('x must reside on CUDA device')
    if f.device != x.device or f.dtype != torch.float32:
        raise RuntimeError('f must be a float32 tensor on the same device as x')
    if x.ndim != 4:
        raise RuntimeError('x must be rank 4')
    if x.numel() == 0:
        raise RuntimeError('x is empty')
    if x.dtype not in (torch.float32, torch.bfloat16, torch.float16):
        raise RuntimeError('x must be float16, bfloat16 or float32')
    if separable:
        fw = fh = f.numel() // 2
    else:
        if f.ndim != 2 or f.shape[0] < 1 or f.shape[1] < 1:
            raise RuntimeError('f must be rank 2 with at least one tap')
        fh, fw = f.shape
    if min(upx, upy, downx, downy) < 1:
        raise RuntimeError('upsampling and downsampling factors must be at least 1')
    n, c, h, w = x.shape
    out_w = (w * upx + padx0 + padx1 - fw + downx) // downx
    out_h = (h * upy + pady0 + pady1 - fh + downy) // downy
    if out_w < 1 or out_h < 1:
        raise RuntimeError('output must be at least 1x1')
    x = x.contiguous()
    f = f.contiguous()
    y = torch.empty([n, c, out_h, out_w], dtype=x.dtype, device=x.device)
    with torch.cuda.device(x.device):
        rc = lib.cno_upfirdn2d(_ptr(x), _ptr(y), _ptr(f), _DT[x.dtype], int(separable), n * c, h, w, out_h, out_w,
                               fw, fh, upx, upy, downx, downy, padx0, pady0, int(bool(flip)), float(gain), _stream(x))
    _lib.check(rc, 'upfirdn2d')
    return y


class _FilteredLReluPlugin:
    @staticmethod
    def filtered_lrelu(x, fu, fd, b, si, up, down, px0, px1, py0, py1, sx, sy, gain, slope, clamp, flip_filter, write_signs,
                       db=None, signs_only=False, affine=None):
        """filtered_lrelu.cpp:8-201 -> (y, so, return_code); return_code -1 = use the generic composition.
        db (extension, optional): zeroed fp32 [C] tensor that receives y.sum([0,2,3]) from the same kernel.
        signs_only (extension, optional): the caller knows `si` was written with clamp = none (no "clamped" codes), which lets
        the kernel skip that test; the default is always correct.
        affine (extension, optional): (a, d) fp32 [N,C] tensors; the op runs on a[n,c] * x + d[n,c] without materialising it
        (CIN -> filtered_lrelu prologue fusion, forward calls only)."""
        if affine is not None:
            a_, d_ = affine
            ok = all(v.dtype == torch.float32 and v.is_contiguous() and v.device == x.device and v.numel() == x.shape[0] * x.shape[1]
                     for v in (a_, d_))
            if not ok or si.numel() > 0:
                raise RuntimeError('affine must be two contiguous float32 [N,C] tensors on the device of x, forward calls only')
        if db is not None and not (db.dtype == torch.float32 and db.is_contiguous() and db.device == x.device and db.numel() == x.shape[1]):
            raise RuntimeError('db must be a contiguous float32 [C] tensor on the device of x')
        lib = _lib.load()
        if not x.is_cuda:
            raise RuntimeError('x must reside on CUDA device')
        if fu.dtype != torch.float32 or fd.dtype != torch.float32:
            raise RuntimeError('fu and fd must be float32')
        if b.dtype != x.dtype:
            raise RuntimeError('x and b must have the same dtype')
        if x.ndim != 4:
            raise RuntimeError('x must be rank 4')
        if x.numel() == 0:
            raise RuntimeError('x is empty')
        if fu.ndim not in (1, 2) or fd.ndim not in (1, 2) or fu.numel() == 0 or fd.numel() == 0:
            raise RuntimeError('fu and fd must be non-empty and rank 1 or 2')
        if b.ndim != 1 or (b.shape[0] != x.shape[1] and b.numel() != 0):      # empty b (extension): no bias, NULL goes to the kernel
            raise RuntimeError('b must be a vector with the same number of channels as x')
        if up < 1 or down < 1:
            raise RuntimeError('up and down must be at least 1')
        none = (torch.empty(0), torch.empty(0), -1)
        if x.dtype not in (torch.float32, torch.bfloat16, torch.float16) or fu.ndim != 1 or fd.ndim != 1:
            return none
        n, c, xh, xw = x.shape
        fut, fdt = fu.shape[0] - 1, fd.shape[0] - 1
        cw = xw * up + (px0 + px1) - fut
        ch = xh * up + (py0 + py1) - fut
        if not (cw > fdt and ch > fdt):
            raise RuntimeError('upsampled buffer must be at least the size of downsampling filter')
        yw = (cw - fdt + (down - 1)) // down
        yh = (ch - fdt + (down - 1)) // down
        if yw < 1 or yh < 1:
            raise RuntimeError('output must be at least 1x1')
        dt = _DT[x.dtype]
        if lib.cno_filtered_lrelu_query(dt, xh, xw, yh, yw, fu.shape[0], fd.shape[0], up, down, px0, py0) != 0:
            return none
        if clamp < 3.0e38 or not (0.0 <= slope <= 1.0):
            return none
        x = x.contiguous()
        y = torch.empty([n, c, yh, yw], dtype=x.dtype, device=x.device)
        read_signs = si.numel() > 0
        so = torch.empty(0)
        mode, s, mask_w = _lib.SIGN_NONE, None, 0
        if write_signs:
            mask_w = yw * down - (down - 1) + fdt
            sh = yh * down - (down - 1) + fdt
            sw = (mask_w + 15) & ~15
            s = so = torch.empty([n, c, sh, sw >> 2], dtype=torch.uint8, device=x.device)
            mode = _lib.SIGN_WRITE
        elif read_signs:
            s = si
            if not (s.is_contiguous() and s.dtype == torch.uint8 and s.device == x.device and s.ndim == 4
                    and s.shape[0] == n and s.shape[1] == c):
                raise RuntimeError('signs must be a contiguous uint8 [N,C,sh,sw/4] tensor on the device of x')
            mask_w = s.shape[3] << 2
            mode = _lib.SIGN_READ_SIGNONLY if signs_only else _lib.SIGN_READ
        tu, td = host_taps(fu), host_taps(fd)
        with torch.cuda.device(x.device):
            rc = lib.cno_filtered_lrelu_affine(_ptr(x), _ptr(y), _ptr(b), _ptr(s), _ptr(db),
                                               tu.ctypes.data_as(ctypes.c_void_p), td.ctypes.data_as(ctypes.c_void_p), dt,
                                               n, c, xh, xw, yh, yw, fu.shape[0], fd.shape[0], up, down, px0, py0,
                                               s.shape[2] if s is not None else 0, mask_w, s.shape[3] if s is not None else 0,
                                               int(sx), int(sy), mode, float(gain), float(slope), -1.0, int(bool(flip_filter)),
                                               _ptr(affine[0]) if affine is not None else None,
                                               _ptr(affine[1]) if affine is not None else None, _stream(x))
        if _lib.check(rc, 'filtered_lrelu') < 0:
            return none
        return y, so, 0

    @staticmethod
    def filtered_lrelu_act_(x, si, sx, sy, gain, slope, clamp, write_signs):
        """filtered_lrelu.cpp:205-282: in place on x, returns the sign tensor when write_signs."""
        lib = _lib.load()
        if not x.is_cuda or x.ndim != 4:
            raise RuntimeError('x must be a rank 4 CUDA tensor')
        if x.dtype not in (torch.float32, torch.bfloat16, torch.float16):
            raise RuntimeError('x must be float16, bfloat16 or float32')
        if not x.is_contiguous():
            raise RuntimeError('x must be contiguous')      # the generic path always produces dense tensors
        n, c, h, w = x.shape
        so = torch.empty(0)
        mode, s, mask_w = _lib.SIGN_NONE, None, 0
        if write_signs:
            s = so = torch.empty([n, c, h, ((w + 15) & ~15) >> 2], dtype=torch.uint8, device=x.device)
            mask_w, mode = w, _lib.SIGN_WRITE
        elif si.numel() > 0:
            s = si
            if not (s.is_contiguous() and s.dtype == torch.uint8 and s.device == x.device and s.ndim == 4
                    and s.shape[0] == n and s.shape[1] == c):
                raise RuntimeError('signs must be a contiguous uint8 [N,C,sh,sw/4] tensor on the device of x')
            mask_w, mode = s.shape[3] << 2, _lib.SIGN_READ
        if x.numel() == 0:
            return so
        with torch.cuda.device(x.device):
            rc = lib.cno_filtered_lrelu_act(_ptr(x), _ptr(s), _DT[x.dtype], n * c, h, w,
                                            s.shape[2] if s is not None else 0, mask_w, s.shape[3] if s is not None else 0,
                                            int(sx), int(sy), mode, float(gain), float(slope),
                                            float(clamp) if clamp < 3.0e38 else -1.0, _stream(x))
        _lib.check(rc, 'filtered_lrelu_act_')
        return so


_PLUGINS = {'bias_act_plugin': _BiasActPlugin, 'upfirdn2d_plugin': _Upfirdn2dPlugin,
            'filtered_lrelu_plugin': _FilteredLReluPlugin}


def get_plugin(module_name, sources=None, headers=None, source_dir=None, **build_kwargs):
    """Same call as the reference (custom_ops.py:53); sources / build flags are ignored because the
    kernels are prebuilt.  Raises if libcno_b200.so is missing."""
    _lib.load()
    try:
        return _PLUGINS[module_name]
    except KeyError:
        raise RuntimeError(f'unknown plugin {module_name!r}; available: {sorted(_PLUGINS)}')
