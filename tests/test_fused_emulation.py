"""CPU emulation of the fused filtered_lrelu kernel (same core header as the CUDA kernel, compiled
with g++) against the oracle: forward + sign mask, and backward through the sign-read mode with the
reference's swapped-op recipe (filtered_lrelu.py:244-258).  No GPU needed."""
import ctypes, math, os, subprocess
import numpy as np
import pytest
import torch
from oracle import cno_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_lib = None


def emu():
    global _lib
    if _lib is None:
        out = os.path.join(ROOT, 'tests', 'emu', 'libflr_emu.so')
        src = os.path.join(ROOT, 'tests', 'emu', 'flr_emu.cpp')
        deps = [src] + [os.path.join(ROOT, 'pregenerating-pde_b200', 'csrc', f) for f in ('flrelu_core.h', 'flrelu_host.h')]
        if not os.path.exists(out) or any(os.path.getmtime(d) > os.path.getmtime(out) for d in deps):
            subprocess.check_call(['g++', '-O1', '-std=c++17', '-shared', '-fPIC', '-o', out, src])
        _lib = ctypes.CDLL(out)
    return _lib


def fp(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def run_emu(x, b, fu, fd, up, down, pad, gain, slope, clamp, flip, mode, mask=None, mask_geom=None, sxy=(0, 0), bands=0, affine=None):
    n, c, h, w = x.shape
    oh, ow = O.filtered_lrelu_out_size((h, w), torch.from_numpy(fu), torch.from_numpy(fd), up, down, pad)
    y = np.full((n, c, oh, ow), np.nan, np.float32)
    if mask_geom is None:
        sh, swb, swv = O.sign_mask_shape((oh, ow), torch.from_numpy(fd), down)
        mask_geom = (sh, swv, swb)
    if mask is None:
        mask = np.zeros((n, c, mask_geom[0], mask_geom[2]), np.uint8)
    info = (ctypes.c_int * 5)()
    bz = b if b is not None else np.zeros(c, np.float32)
    pa, pd = (fp(affine[0]), fp(affine[1])) if affine is not None else (None, None)
    rc = emu().flr_emulate_affine(fp(x), fp(y), fp(bz) if b is not None else None, fp(mask), n, c, h, w, oh, ow,
                                  fp(fu), len(fu), fp(fd), len(fd), up, down, pad[0], pad[2],
                                  mask_geom[0], mask_geom[1], mask_geom[2], sxy[0], sxy[1], mode,
                                  ctypes.c_float(gain), ctypes.c_float(slope), ctypes.c_float(-1.0 if clamp is None else clamp),
                                  int(flip), bands, info, pa, pd)
    return rc, y, mask, list(info)


def rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


CASES = [  # name, in, out, N, C, forced number of persistent CTAs (0 = heuristic)
    ('I', 16, 16, 2, 3, 0), ('D', 16, 8, 2, 3, 0), ('U', 8, 16, 2, 3, 0),
    ('I', 32, 32, 1, 5, 0), ('D', 32, 16, 1, 2, 0), ('U', 16, 32, 1, 2, 0),
    ('I', 64, 64, 1, 2, 2), ('D', 64, 32, 1, 2, 3), ('U', 32, 64, 1, 3, 5), ('I', 128, 128, 1, 1, 4),
    ('I', 8, 8, 3, 40, 0), ('I', 32, 32, 2, 9, 7), ('U', 16, 32, 2, 9, 11), ('D', 32, 16, 3, 7, 6),
    # planes wider than one 256-element TMA box (BASELINE.json configs[4], 256^2 fine-tune): several overlapping boxes per strip
    ('I', 256, 256, 1, 2, 3), ('D', 256, 128, 1, 1, 2), ('U', 128, 256, 1, 1, 2), ('I', 512, 512, 1, 1, 5),
]


@pytest.mark.parametrize('name,s_in,s_out,N,C,bands', CASES)
def test_emulated_forward_backward(name, s_in, s_out, N, C, bands):
    cfg = O.lrelu_config(s_in, s_out)
    fu, fd, up, down, pad = cfg['fu'], cfg['fd'], cfg['up'], cfg['down'], cfg['padding']
    g = torch.Generator().manual_seed(s_in * 7 + s_out)
    x = torch.randn(N, C, s_in, s_in, generator=g)
    b = torch.randn(C, generator=g)
    gain, slope = math.sqrt(2), 0.2
    xr = x.clone().requires_grad_(True); br = b.clone().requires_grad_(True)
    y_ref = O.filtered_lrelu(xr, fu, fd, br, up=up, down=down, padding=pad, gain=gain, slope=slope)
    dy = torch.randn(y_ref.shape, generator=g)
    dx_ref, db_ref = torch.autograd.grad(y_ref, [xr, br], dy)

    rc, y, mask, info = run_emu(x.numpy(), b.numpy(), fu.numpy(), fd.numpy(), up, down, pad, gain, slope, None, False, 1, bands=bands)
    assert rc == 0, 'fused plan refused a CNO configuration'
    assert not np.isnan(y).any()
    assert rel(y, y_ref.detach().numpy()) < 2e-6
    # sign mask == sign of the oracle pre-activation (ambiguous only where |v| is at round-off)
    v = O.filtered_lrelu_preact(x, fu, b, up=up, padding=pad, gain=gain)
    sh, swb, swv = O.sign_mask_shape(y.shape[2:], fd, down)
    code_ref = O.sign_mask_from_preact(v, slope, gain, None, sh, swv)
    code = O.unpack_sign_mask(torch.from_numpy(mask), swv)
    diff = code != code_ref
    assert int((diff & (v[:, :, :sh, :swv].abs() > 1e-5)).sum()) == 0
    assert int(diff.sum()) <= 2

    # backward = same op, up<->down, fu<->fd, flipped, mask read with offsets (filtered_lrelu.py:244-256)
    tu, td = len(fu), len(fd)
    pp = [(tu - 1) + (td - 1) - pad[0], s_in * up - s_out * down + pad[0] - (up - 1),
          (tu - 1) + (td - 1) - pad[2], s_in * up - s_out * down + pad[2] - (up - 1)]
    gg = gain * up ** 2 / down ** 2
    sx, sy = pad[0] - (tu - 1), pad[2] - (tu - 1)
    rc, dx, _, _ = run_emu(dy.numpy(), None, fd.numpy(), fu.numpy(), down, up, pp, gg, slope, None, True, 2,
                           mask=mask, mask_geom=(sh, swv, swb), sxy=(sx, sy), bands=bands)
    assert rc == 0
    assert dx.shape == tuple(x.shape)
    assert rel(dx, dx_ref.numpy()) < 2e-6
    assert rel(dx.sum((0, 2, 3)), db_ref.numpy()) < 1e-5
    # sign-only read mode (mask known to be written with clamp = none): identical result
    rc, dx_so, _, _ = run_emu(dy.numpy(), None, fd.numpy(), fu.numpy(), down, up, pp, gg, slope, None, True, 3,
                              mask=mask, mask_geom=(sh, swv, swb), sxy=(sx, sy), bands=bands)
    assert rc == 0 and np.array_equal(dx_so, dx)


def test_emulated_flip_asymmetric_none_mode():
    g = torch.Generator().manual_seed(3)
    fu = torch.randn(12, generator=g); fd = torch.randn(12, generator=g)
    x = torch.randn(2, 3, 12, 12, generator=g); b = torch.randn(3, generator=g)
    pad = [11, 10, 11, 10]
    for flip in (True, False):
        y_ref = O.filtered_lrelu(x, fu, fd, b, up=2, down=2, padding=pad, gain=1.3, slope=0.1, flip_filter=flip)
        rc, y, mask, _ = run_emu(x.numpy(), b.numpy(), fu.numpy(), fd.numpy(), 2, 2, pad, 1.3, 0.1, None, flip, 1)
        assert rc == 0 and rel(y, y_ref.numpy()) < 2e-6
        v = O.filtered_lrelu_preact(x, fu, b, up=2, padding=pad, gain=1.3, flip_filter=flip)
        sh, swb, swv = O.sign_mask_shape(y.shape[2:], fd, 2)
        code_ref = O.sign_mask_from_preact(v, 0.1, 1.3, None, sh, swv)
        assert int((O.unpack_sign_mask(torch.from_numpy(mask), swv) != code_ref).sum()) <= 2
        # pad bytes of every mask row are written as zeros
        assert int(torch.from_numpy(mask)[..., (swv + 3) // 4:].sum()) == 0
        rc, y0, _, _ = run_emu(x.numpy(), b.numpy(), fu.numpy(), fd.numpy(), 2, 2, pad, 1.3, 0.1, None, flip, 0)
        assert rc == 0 and np.array_equal(y0, y)


@pytest.mark.parametrize('pad', [[10, 11, 10, 11], [12, 9, 12, 9]])
def test_emulated_other_padding_phase(pad):
    """px0 even -> the other polyphase variant (window of 7 input samples)."""
    g = torch.Generator().manual_seed(4)
    cfg = O.lrelu_config(16, 16)
    x = torch.randn(1, 2, 16, 16, generator=g); b = torch.randn(2, generator=g)
    y_ref = O.filtered_lrelu(x, cfg['fu'], cfg['fd'], b, up=2, down=2, padding=pad)
    rc, y, _, _ = run_emu(x.numpy(), b.numpy(), cfg['fu'].numpy(), cfg['fd'].numpy(), 2, 2, pad, math.sqrt(2), 0.2, None, False, 1)
    assert rc == 0 and y.shape == tuple(y_ref.shape) and rel(y, y_ref.numpy()) < 2e-6


def test_plan_refuses_unsupported():
    x = np.zeros((1, 1, 8, 8), np.float32); f = np.ones(4, np.float32) / 4
    rc, *_ = run_emu(x, None, f, f, 2, 2, [1, 1, 1, 1], 1.0, 0.2, None, False, 0)
    assert rc == -1   # 4-tap filters have no fused variant -> generic composition
    cfg = O.lrelu_config(8, 8)
    x = np.zeros((1, 1, 8, 8), np.float32)
    for kw in (dict(clamp=0.5, slope=0.2), dict(clamp=None, slope=1.5)):   # clamp / slope > 1 -> generic composition
        rc, *_ = run_emu(x, None, cfg['fu'].numpy(), cfg['fd'].numpy(), 2, 2, cfg['padding'], 1.0, kw['slope'], kw['clamp'], False, 0)
        assert rc == -1


@pytest.mark.parametrize('name,s_in,s_out,N,C,bands', [CASES[0], CASES[1], CASES[2], CASES[6], CASES[12]])
def test_emulated_affine_prologue(name, s_in, s_out, N, C, bands):
    """AFF variants: the kernel fed x with the plane affine (a, d) equals the plain kernel math on a*x + d --
    checked against the oracle's filtered_lrelu of the materialised tensor, outputs and sign mask."""
    cfg = O.lrelu_config(s_in, s_out)
    fu, fd, up, down, pad = cfg['fu'], cfg['fd'], cfg['up'], cfg['down'], cfg['padding']
    g = torch.Generator().manual_seed(s_in * 11 + s_out)
    x = torch.randn(N, C, s_in, s_in, generator=g) * 3 + 1
    b = torch.randn(C, generator=g)
    a = torch.randn(N, C, generator=g)
    d = torch.randn(N, C, generator=g)
    gain, slope = math.sqrt(2), 0.2
    z = x * a[:, :, None, None] + d[:, :, None, None]
    y_ref = O.filtered_lrelu(z, fu, fd, b, up=up, down=down, padding=pad, gain=gain, slope=slope)
    for mode in (0, 1):
        rc, y, mask, _ = run_emu(x.numpy(), b.numpy(), fu.numpy(), fd.numpy(), up, down, pad, gain, slope, None, False, mode, bands=bands,
                                 affine=(a.numpy().copy(), d.numpy().copy()))
        assert rc == 0
        assert not np.isnan(y).any()
        assert rel(y, y_ref.numpy()) < 2e-5          # fp32, different summation order (a * FIR(x) + d * edge vs FIR(a*x + d))
    # the mask written by the AFF variant drives the plain backward to the gradient of the materialised composition
    rc, _, mask0, _ = run_emu(z.numpy(), b.numpy(), fu.numpy(), fd.numpy(), up, down, pad, gain, slope, None, False, 1, bands=bands)
    differ = np.unpackbits(mask ^ mask0).sum()
    assert differ <= max(2, mask.size // 20000)      # sign flips only where the pre-activation is ~0 (rounding)
