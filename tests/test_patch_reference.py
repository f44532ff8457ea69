"""patch_reference(): the zero-edit route of INTEGRATION.md §1.

CPU (build container only, skipped where /root/reference is absent): the UNMODIFIED reference modules are imported in a
subprocess, patched, and a time-conditioned CNOBlock / ResidualBlock must give the same output and gradients as before the
patch (CPU tensors take the modules' own torch ops through our replaced forwards), and unpatch must restore every attribute.

GPU (reference absent on the box): a stand-in module tree with the reference's names and call shapes (torch_utils.custom_ops,
torch_utils.ops.*, CNO_timeModule_CIN with eager FILM / CNOBlock / ResidualBlock written as CNO_timeModule_CIN.py:60-70,
:329-335, :496-510) is patched; the patched blocks must run the fused kernels and agree with the stand-in's eager result."""
import os
import subprocess
import sys
import textwrap
import types

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = '/root/reference/CNO_Experiments'

_CPU_SCRIPT = r'''
import sys, types, copy, torch
sys.path.insert(0, ROOT)
pl = types.ModuleType('pytorch_lightning')
class LightningModule(torch.nn.Module):
    def log(self, *a, **k): pass
    def save_hyperparameters(self, *a, **k): pass
pl.LightningModule = LightningModule
sys.modules['pytorch_lightning'] = pl
for name in ['matplotlib', 'matplotlib.pyplot', 'h5py', 'netCDF4']:
    sys.modules[name] = types.ModuleType(name)
sys.modules['matplotlib'].pyplot = sys.modules['matplotlib.pyplot']
sys.path.insert(0, REF)
import CNO_timeModule_CIN as R
import torch_utils.custom_ops as r_ops
import pregenerating_pde_b200 as cno

def run(block, x, t):
    xr = x.clone().requires_grad_(True)
    y = block(xr, t)
    ps = [xr] + list(block.parameters())
    g = torch.autograd.grad(y, ps, torch.ones_like(y) * 0.37, allow_unused=True)      # unused: a convolution bias in front of a CIN
    return [y] + [torch.zeros_like(p) if gi is None else gi for p, gi in zip(ps, g)]  # (cin.conv_before_film), reference gradient = noise

torch.manual_seed(3)
kw = dict(is_time=True, nl_dim=[2, 3], batch_norm=False)
blocks = [R.CNOBlock(3, 4, 16, 8, **kw), R.CNOBlock(4, 4, 8, 16, **kw), R.ResidualBlock(4, 16, **kw)]
for b in blocks:
    with torch.no_grad():
        for p in b.parameters():
            p.add_(0.2 * torch.randn_like(p))
xs = [torch.randn(2, 3, 16, 16), torch.randn(2, 4, 8, 8), torch.randn(2, 4, 16, 16)]
t = torch.rand(2)
before = [run(b, x, t) for b, x in zip(blocks, xs)]
orig = (r_ops.get_plugin, R.FILM.forward, R.CNOBlock.forward, R.ResidualBlock.forward)
assert cno.patch_reference(REF)
assert r_ops.get_plugin is not orig[0] and R.FILM.forward is cno.cin.film_forward
assert R.CNOBlock.forward is not orig[2] and R.ResidualBlock.forward is not orig[3]
after = [run(b, x, t) for b, x in zip(blocks, xs)]
for bs, as_ in zip(before, after):
    # a convolution bias in front of an instance norm has a zero gradient (round-off noise in the reference, exactly zero after
    # the patch): errors are measured against the block's gradient scale, not against each tensor's own maximum
    floor = 1e-3 * max(float(u.detach().abs().max()) for u in bs[1:])
    for u, v in zip(bs, as_):
        assert u.shape == v.shape
        err = float((u - v).abs().max()) / max(float(u.abs().max()), floor)
        assert err < 1e-4, err      # 1e-7 rounding differences of the bias-free convolution + leaky-ReLU kinks (tests/test_cno_model.py)
# flavours the patch does not cover run the reference's own forward
bn = R.CNOBlock(3, 4, 16, 16, is_time=False, nl_dim=[2, 3], batch_norm=True)
y = bn(xs[0], t)
assert y.shape == (2, 4, 16, 16)
cno.unpatch_reference()
assert (r_ops.get_plugin, R.FILM.forward, R.CNOBlock.forward, R.ResidualBlock.forward) == orig
print('PATCH_OK')
'''


@pytest.mark.skipif(not os.path.isdir(REF), reason='reference checkout not present (GPU box)')
def test_patch_reference_cpu_roundtrip():
    script = f'ROOT = {ROOT!r}\nREF = {REF!r}\n' + _CPU_SCRIPT
    r = subprocess.run([sys.executable, '-c', script], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and 'PATCH_OK' in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def _standin_modules():
    """module tree with the reference's names; bodies restate CNO_timeModule_CIN.py:60-70, :329-335, :496-510 in eager torch"""
    from pregenerating_pde_b200 import cno_model
    import torch.nn as nn
    mods = {}
    for name in ['torch_utils', 'torch_utils.ops', 'torch_utils.custom_ops', 'torch_utils.ops.bias_act', 'torch_utils.ops.upfirdn2d',
                 'torch_utils.ops.filtered_lrelu', 'CNO_timeModule_CIN']:
        mods[name] = types.ModuleType(name)
    mods['torch_utils'].__path__ = []
    mods['torch_utils.ops'].__path__ = []

    def no_plugin(*a, **k):
        raise RuntimeError('stand-in: the reference would JIT-compile its plugin here')
    mods['torch_utils.custom_ops'].get_plugin = no_plugin
    for n in ('bias_act', 'upfirdn2d', 'filtered_lrelu'):
        mods[f'torch_utils.ops.{n}']._plugin = 'stale'

    class FILM(nn.Module):
        def __init__(self, channels, s=16, intermediate=128):
            super().__init__()
            self.inp2lat_sacale = nn.Linear(1, intermediate); self.lat2scale = nn.Linear(intermediate, channels)
            self.inp2lat_bias = nn.Linear(1, intermediate); self.lat2bias = nn.Linear(intermediate, channels)
            self.norm = nn.InstanceNorm2d(channels, affine=True)

        def forward(self, x, timestep):
            timestep = timestep.reshape(-1, 1).type_as(x)
            scale = self.lat2scale(self.inp2lat_sacale(timestep))
            bias = self.lat2bias(self.inp2lat_bias(timestep))
            return self.norm(x) * scale.unsqueeze(2).unsqueeze(3) + bias.unsqueeze(2).unsqueeze(3)

    class LReLu(cno_model.LReLu):                 # same buffers / attributes / call as training/filtered_networks.py:200-278
        pass

    def act(c, s_in, s_out):
        return LReLu(c, c, s_in, s_out, **cno_model._filter_args(s_in, s_out, 2.0001, 0.8))

    class CNOBlock(nn.Module):
        def __init__(self, cin, cout, s_in, s_out, is_time=True):
            super().__init__()
            self.convolution = nn.Conv2d(cin, cout, 3, padding=1)
            self.batch_norm = nn.Identity()
            self.activation = act(cout, s_in, s_out)
            self.is_time = is_time
            if is_time:
                self.in_norm_conditiner = FILM(cout, s=s_in)

        def forward(self, x, time):
            x = self.batch_norm(self.convolution(x))
            if self.is_time == 1 or self.is_time == True:
                x = self.in_norm_conditiner(x, time)
            return self.activation(x)

    class ResidualBlock(nn.Module):
        def __init__(self, c, s, is_time=True):
            super().__init__()
            self.convolution1, self.convolution2 = nn.Conv2d(c, c, 3, padding=1), nn.Conv2d(c, c, 3, padding=1)
            self.batch_norm1, self.batch_norm2 = nn.Identity(), nn.Identity()
            self.activation = act(c, s, s)
            self.is_time = is_time
            if is_time:
                self.in_norm_conditiner1, self.in_norm_conditiner2 = FILM(c, s=s), FILM(c, s=s)

        def forward(self, x, time):
            out = self.batch_norm1(self.convolution1(x))
            if self.is_time == 1 or self.is_time == True:
                out = self.in_norm_conditiner1(out, time)
            out = self.batch_norm2(self.convolution2(self.activation(out)))
            if self.is_time == 1 or self.is_time == True:
                out = self.in_norm_conditiner2(out, time)
            return x + out

    m = mods['CNO_timeModule_CIN']
    m.FILM, m.LReLu, m.CNOBlock, m.ResidualBlock = FILM, LReLu, CNOBlock, ResidualBlock
    return mods


@pytest.mark.gpu
def test_patch_reference_standin_gpu():
    import pregenerating_pde_b200 as cno
    from pregenerating_pde_b200 import cno_model
    assert torch.cuda.is_available()
    clash = [n for n in ('torch_utils', 'CNO_timeModule_CIN') if n in sys.modules]
    assert not clash, f'real reference modules already imported: {clash}'
    mods = _standin_modules()
    sys.modules.update(mods)
    R = mods['CNO_timeModule_CIN']
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False                # convolutions on both sides in full fp32
    try:
        torch.manual_seed(5)
        blocks = [R.CNOBlock(3, 8, 32, 32), R.CNOBlock(4, 6, 32, 16), R.CNOBlock(4, 6, 16, 32), R.ResidualBlock(8, 32), R.CNOBlock(3, 4, 16, 16, is_time=False)]
        xs = [torch.randn(3, 3, 32, 32), torch.randn(3, 4, 32, 32), torch.randn(3, 4, 16, 16), torch.randn(3, 8, 32, 32), torch.randn(3, 3, 16, 16)]
        t = torch.rand(3, device='cuda')
        for b in blocks:
            with torch.no_grad():
                for p in b.parameters():
                    p.add_(0.2 * torch.randn_like(p))
            b.cuda()
            for a in b.modules():
                if isinstance(a, cno_model.LReLu):
                    a.impl = 'ref'                          # the unpatched stand-in is pure torch ops
        xs = [x.cuda() for x in xs]

        def run(b, x):
            xr = x.clone().requires_grad_(True)
            y = b(xr, t)
            dy = torch.linspace(-1, 1, y.numel(), device='cuda').reshape(y.shape)
            ps = [xr] + list(b.parameters())
            gs = torch.autograd.grad(y, ps, dy, allow_unused=True)       # unused: a convolution bias in front of a CIN (cin.conv_before_film)
            return y, [y] + [torch.zeros_like(p) if gi is None else gi for p, gi in zip(ps, gs)]
        before = [run(b, x)[1] for b, x in zip(blocks, xs)]
        assert cno.patch_reference()
        for n in ('bias_act', 'upfirdn2d', 'filtered_lrelu'):
            assert mods[f'torch_utils.ops.{n}']._plugin is None          # forces re-resolution through our loader
        assert mods['torch_utils.custom_ops'].get_plugin('filtered_lrelu_plugin') is cno.custom_ops.get_plugin('filtered_lrelu_plugin')
        assert R.FILM.forward is cno.cin.film_forward
        for b in blocks:
            for a in b.modules():
                if isinstance(a, cno_model.LReLu):
                    a.impl = 'cuda'
        after, fns = [], []
        for b, x in zip(blocks, xs):
            y, outs = run(b, x)
            after.append(outs)
            names, node, seen = set(), [y.grad_fn], set()
            while node:                                    # autograd graph of the patched block
                f = node.pop()
                if f is None or f in seen:
                    continue
                seen.add(f); names.add(type(f).__name__)
                node += [nf for nf, _ in f.next_functions]
            fns.append(names)
        assert all(any('FilmLRelu' in n for n in s) for s in fns[:4]), fns       # CIN folded into the activation
        assert any(n.startswith('_Film') and 'LRelu' not in n for n in fns[3])    # ResidualBlock: second CIN with the residual epilogue
        assert not any('Film' in n for n in fns[4])                               # not time-conditioned: the stand-in's own forward
        for bi, (bs, as_) in enumerate(zip(before, after)):
            # a convolution bias in front of an instance norm has a zero gradient (rounding noise on both sides): errors are
            # measured against the block's gradient scale, not against each tensor's own maximum
            floor = 1e-3 * max(float(u.detach().abs().max()) for u in bs[1:])
            for ti, (u, v) in enumerate(zip(bs, as_)):
                err = float((u - v).detach().abs().max()) / max(float(u.detach().abs().max()), floor)
                assert u.shape == v.shape and err < 5e-4, (bi, ti, err)
        cno.unpatch_reference()
        assert R.FILM.forward is not cno.cin.film_forward
        assert mods['torch_utils.custom_ops'].get_plugin.__name__ == 'no_plugin'
    finally:
        torch.backends.cudnn.allow_tf32 = tf32
        cno.unpatch_reference()
        for n in mods:
            sys.modules.pop(n, None)
