"""Second, GPU-side oracle: the reference's OWN CUDA plugins (torch_utils/ops/*.cu, unmodified sources, compiled for
sm_100a by oracle/build_ref_plugins.py into oracle/_ref/plugins/) run next to our kernels on the same device buffers.
Outputs must agree to fp32 round-off and the 2-bit sign mask BYTES -- layout, padding bytes and codes -- must be identical
wherever the pre-activation is not at rounding distance from zero.  Skipped when the plugins were not built."""
import math
import os
import sys

import pytest
import torch

from oracle import cno_oracle as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEV = 'cuda'


@pytest.fixture(scope='module')
def ref():
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import build_ref_plugins as B
    mods = {n: B.load(n) for n in B.PLUGINS}
    if any(m is None for m in mods.values()):
        pytest.skip('oracle/_ref/plugins not built (python oracle/build_ref_plugins.py in the build container)')
    return mods


@pytest.fixture(scope='module')
def cno():
    import pregenerating_pde_b200 as pkg
    pkg._lib.load()
    return pkg


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


CASES = [('I', 128, 128, 2, 4), ('D', 128, 64, 2, 3), ('U', 64, 128, 2, 3), ('I', 32, 32, 2, 8), ('I', 8, 8, 3, 16), ('I', 256, 256, 1, 2)]


@pytest.mark.parametrize('dtype', [torch.float32, torch.float16])
@pytest.mark.parametrize('name,s_in,s_out,N,C', CASES)
def test_filtered_lrelu_vs_reference_cuda_kernel(cno, ref, name, s_in, s_out, N, C, dtype):
    cfg = O.lrelu_config(s_in, s_out)
    up, down = cfg['up'], cfg['down']
    px0, px1, py0, py1 = cfg['padding']
    g = torch.Generator().manual_seed(s_in + 5 * s_out + C)
    x = torch.randn(N, C, s_in, s_in, generator=g).to(DEV, dtype)
    b = torch.randn(C, generator=g).to(DEV, dtype)
    dy = torch.randn(N, C, s_out, s_out, generator=g).to(DEV, dtype)
    fu, fd = cfg['fu'].to(DEV), cfg['fd'].to(DEV)
    tu, td = fu.numel(), fd.numel()
    gain, slope, inf, e = math.sqrt(2), 0.2, float('inf'), torch.empty(0)
    ours = cno.custom_ops.get_plugin('filtered_lrelu_plugin')
    theirs = ref['filtered_lrelu_plugin']
    args = (x, fu, fd, b, e, up, down, px0, px1, py0, py1, 0, 0, gain, slope, inf, False, True)
    y_o, so_o, rc_o = ours.filtered_lrelu(*args)
    y_r, so_r, rc_r = theirs.filtered_lrelu(*args)
    assert rc_o == 0
    if rc_r != 0:
        pytest.skip('the reference has no specialised kernel for this call')
    tol = 1e-5 if dtype == torch.float32 else 2e-3
    assert rel(y_o, y_r) < tol
    assert so_o.shape == so_r.shape and so_o.dtype == so_r.dtype
    # Codes of the real up-res columns (x < UW) must agree; they can only differ where |pre-activation| is at fp32 rounding
    # distance from 0 (different summation orders).  The reference pads mask rows to 16 elements (filtered_lrelu.cpp:83-85) and
    # its tiles write whatever sign they computed into those pad columns; ours writes zeros there.  No reader ever uses them
    # (the backward kernels only index columns < UW), so pad bits are excluded from the comparison.
    uw = s_out * down - (down - 1) + (td - 1)
    codes = lambda m: torch.stack([(m >> (2 * i)) & 3 for i in range(4)], dim=-1).flatten(-2)[..., :uw]
    differ = int((codes(so_o) != codes(so_r)).sum())
    assert differ <= max(2, so_o.numel() // 50000), f'{differ} of {so_o.numel() * 4} codes differ'
    # backward: each kernel reads the OTHER implementation's mask (layout interchange both ways)
    pp = (tu + td - 2 - px0, s_in * up - s_out * down + px0 - (up - 1), tu + td - 2 - py0, s_in * up - s_out * down + py0 - (up - 1))
    gg = gain * up * up / (down * down)
    zb = torch.zeros_like(b)
    bargs = lambda m: (dy, fd, fu, zb, m, down, up, pp[0], pp[1], pp[2], pp[3], px0 - (tu - 1), py0 - (tu - 1), gg, slope, inf, True, False)
    dx_o, _, rc1 = ours.filtered_lrelu(*bargs(so_r))
    dx_r, _, rc2 = theirs.filtered_lrelu(*bargs(so_o))
    assert rc1 == 0
    if rc2 == 0:
        assert rel(dx_o, dx_r) < (1e-5 if dtype == torch.float32 else 2e-3) or differ > 0


def test_upfirdn2d_and_bias_act_vs_reference_cuda_kernels(cno, ref):
    g = torch.Generator().manual_seed(3)
    x = torch.randn(2, 5, 64, 64, generator=g).to(DEV)
    f = O.lrelu_config(64, 64)['fu'].to(DEV)
    f2 = torch.outer(f, f).contiguous()
    for (u, d, pad) in [(2, 1, (6, 5, 6, 5)), (1, 2, (5, 5, 5, 5)), (1, 1, (6, 5, 6, 5))]:
        y_r = ref['upfirdn2d_plugin'].upfirdn2d(x, f2, u, u, d, d, pad[0], pad[1], pad[2], pad[3], False, float(u * u))
        y_o = cno.custom_ops.get_plugin('upfirdn2d_plugin').upfirdn2d(x, f2, u, u, d, d, pad[0], pad[1], pad[2], pad[3], False, float(u * u))
        assert y_o.shape == y_r.shape and rel(y_o, y_r) < 1e-5
    b = torch.randn(5, generator=g).to(DEV)
    e = torch.empty(0)
    for act in range(1, 10):
        a_r = ref['bias_act_plugin'].bias_act(x, b, e, e, e, 0, 1, act, 0.2, 1.3, -1.0)
        a_o = cno.custom_ops.get_plugin('bias_act_plugin').bias_act(x, b, e, e, e, 0, 1, act, 0.2, 1.3, -1.0)
        assert rel(a_o, a_r) < 1e-5          # the reference builds with --use_fast_math (bias_act.py:35)
