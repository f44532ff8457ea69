"""Pins oracle/cno_oracle.py against vectors produced by the unmodified reference
(oracle/make_golden.py -> tests/golden/*.npz).  CPU only."""
import math
import numpy as np
import torch
import pytest
from oracle import cno_oracle as O

T = lambda a: torch.from_numpy(np.asarray(a))


def relerr(a, b):
    a, b = a.detach().double(), b.detach().double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


FL_CASES = ['I16', 'D16', 'U8', 'I32', 'D32', 'U16', 'I8']


@pytest.mark.parametrize('name', FL_CASES)
def test_filtered_lrelu_matches_reference(golden, name):
    g = golden['filtered_lrelu']
    up, down, *pad = [int(v) for v in g[f'{name}.cfg']]
    x = T(g[f'{name}.x']).requires_grad_(True)
    b = T(g[f'{name}.b']).requires_grad_(True)
    y = O.filtered_lrelu(x, T(g[f'{name}.fu']), T(g[f'{name}.fd']), b, up=up, down=down, padding=pad)
    dx, db = torch.autograd.grad(y, [x, b], T(g[f'{name}.dy']))
    assert relerr(y, T(g[f'{name}.y'])) < 1e-6
    assert relerr(dx, T(g[f'{name}.dx'])) < 1e-6
    assert relerr(db, T(g[f'{name}.db'])) < 1e-6
    v = O.filtered_lrelu_preact(x.detach(), T(g[f'{name}.fu']), b.detach(), up=up, padding=pad)
    assert relerr(v, T(g[f'{name}.v'])) < 1e-6
    # LReLu's filter design / padding formula restated in the oracle
    cfg = O.lrelu_config(x.shape[-1], y.shape[-1])
    assert (cfg['up'], cfg['down'], cfg['padding']) == (up, down, pad)
    assert torch.allclose(cfg['fu'], T(g[f'{name}.fu'])) and torch.allclose(cfg['fd'], T(g[f'{name}.fd']))


def test_filtered_lrelu_generic_args(golden):
    g = golden['filtered_lrelu']
    x = T(g['GEN.x']).requires_grad_(True)
    b = T(g['GEN.b']).requires_grad_(True)
    y = O.filtered_lrelu(x, T(g['GEN.fu']), T(g['GEN.fd']), b, up=2, down=2, padding=[5, 4, 6, 3], gain=1.3,
                         slope=0.1, clamp=0.7, flip_filter=True)
    dx, db = torch.autograd.grad(y, [x, b], T(g['GEN.dy']))
    assert relerr(y, T(g['GEN.y'])) < 1e-6 and relerr(dx, T(g['GEN.dx'])) < 1e-6 and relerr(db, T(g['GEN.db'])) < 1e-6


UP_SPECS = {'up2': dict(up=2, down=1, padding=[6, 5, 6, 5], gain=4),
            'down2': dict(up=1, down=2, padding=[5, 5, 5, 5], gain=1),
            'xy': dict(up=[2, 1], down=[1, 3], padding=[3, 2, 4, 1], gain=1.5, flip_filter=True),
            'f2d': dict(up=2, down=2, padding=[2, 1, 0, 3], gain=1),
            'crop': dict(up=1, down=1, padding=[-1, -2, 1, -1], gain=1)}


@pytest.mark.parametrize('name', list(UP_SPECS))
def test_upfirdn2d_matches_reference(golden, name):
    g = golden['upfirdn2d']
    x = T(g[f'{name}.x']).requires_grad_(True)
    y = O.upfirdn2d(x, T(g[f'{name}.f']), **UP_SPECS[name])
    dx, = torch.autograd.grad(y, [x], T(g[f'{name}.dy']))
    assert y.shape == g[f'{name}.y'].shape
    assert relerr(y, T(g[f'{name}.y'])) < 1e-6 and relerr(dx, T(g[f'{name}.dx'])) < 1e-6


@pytest.mark.parametrize('act', list(O.ACTIVATIONS))
def test_bias_act_matches_reference(golden, act):
    g = golden['bias_act']
    for tag, kw in [('def', {}), ('arg', dict(alpha=0.3, gain=1.7, clamp=0.9))]:
        x = T(g['x']).requires_grad_(True); b = T(g['b']).requires_grad_(True)
        y = O.bias_act(x, b, act=act, **kw)
        dx, db = torch.autograd.grad(y, [x, b], T(g['dy']))
        assert relerr(y, T(g[f'{act}.{tag}.y'])) < 1e-6
        assert relerr(dx, T(g[f'{act}.{tag}.dx'])) < 1e-6
        assert relerr(db, T(g[f'{act}.{tag}.db'])) < 1e-6


@pytest.mark.parametrize('name', ['c4s16', 'c6s8'])
def test_film_matches_reference(golden, name):
    g = golden['cin']
    w = {k[len(name) + 3:]: T(g[k]).requires_grad_(True) for k in g.files if k.startswith(name + '.p.')}
    x = T(g[f'{name}.x']).requires_grad_(True)
    y = O.film_forward(x, T(g[f'{name}.t']), w)
    names = sorted(w)
    grads = torch.autograd.grad(y, [x] + [w[n] for n in names], T(g[f'{name}.dy']))
    assert relerr(y, T(g[f'{name}.y'])) < 1e-5
    assert relerr(grads[0], T(g[f'{name}.dx'])) < 1e-5
    for n, gr in zip(names, grads[1:]):
        assert relerr(gr, T(g[f'{name}.g.{n}'])) < 2e-5, n


def test_sign_mask_pack_roundtrip_and_shape():
    # C1 numbers from BASELINE.md: sh=266, 68 bytes per row
    fd = torch.zeros(12)
    assert O.sign_mask_shape((128, 128), fd, 2) == (266, 68, 266)
    code = torch.randint(0, 3, (2, 3, 10, 21), dtype=torch.uint8)
    packed = O.pack_sign_mask(code, 8)
    assert packed.shape == (2, 3, 10, 8)
    assert torch.equal(O.unpack_sign_mask(packed, 21), code)


def test_identities():
    # upfirdn2d(x, None) is the identity; filtered_lrelu(up=down=1) == bias_act(lrelu)
    x = torch.randn(2, 3, 5, 7); b = torch.randn(3)
    assert torch.equal(O.upfirdn2d(x, None), x)
    assert torch.allclose(O.filtered_lrelu(x, b=b), O.bias_act(x, b, act='lrelu'))
