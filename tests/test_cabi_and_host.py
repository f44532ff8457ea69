"""CPU checks of the drop-in boundary: the C-ABI library loads, exports exactly what include/cno_b200.h
declares with the argument counts the ctypes binding uses, the Python op wrappers mirror the reference
signatures, and impl='ref' of the package agrees with the oracle.  No compute calls need a GPU."""
import ctypes, inspect, os, re
import numpy as np
import pytest
import torch

from oracle import cno_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope='module')
def pkg():
    import pregenerating_pde_b200 as p
    if not os.path.exists(p._lib.LIB_PATH):
        import importlib.util as u
        spec = u.spec_from_file_location('cno_build', os.path.join(ROOT, 'pregenerating-pde_b200', 'build.py'))
        m = u.module_from_spec(spec); spec.loader.exec_module(m); m.build(verbose=False)
    return p


def header_decls():
    src = open(os.path.join(ROOT, 'include', 'cno_b200.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    out = {}
    for m in re.finditer(r'\b(?:int|const char\*)\s+(cno_\w+)\s*\(([^;]*?)\)\s*;', src, flags=re.S):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if args in ('', 'void') else len(args.split(','))
    return out


def test_library_exports_every_declared_symbol(pkg):
    decls = header_decls()
    assert len(decls) >= 10
    lib = pkg._lib.load()
    for name, nargs in decls.items():
        assert hasattr(lib, name), f'{name} declared in cno_b200.h but not exported'
        assert name in pkg._lib.SIGNATURES, f'{name} has no ctypes signature'
        assert len(pkg._lib.SIGNATURES[name][1]) == nargs, f'{name}: binding has {len(pkg._lib.SIGNATURES[name][1])} args, header {nargs}'
    assert lib.cno_version() >= 100
    assert lib.cno_last_error() is not None


def test_query_mirrors_kernel_coverage(pkg):
    q = pkg._lib.load().cno_filtered_lrelu_query
    # the three CNO configurations at every U-Net level, fp32 and bf16
    for dt in (0, 1):
        for s in (128, 64, 32, 16, 8):
            assert q(dt, s, s, s, s, 12, 12, 2, 2, 11, 11) == 0
            if s >= 16:
                assert q(dt, s, s, s // 2, s // 2, 12, 24, 2, 4, 16, 16) == 0
            if s <= 64:
                assert q(dt, s, s, 2 * s, 2 * s, 24, 12, 4, 2, 18, 18) == 0
    assert q(0, 64, 64, 64, 64, 8, 8, 2, 2, 7, 7) == -1       # other tap counts: generic composition
    assert q(0, 64, 64, 64, 64, 1, 1, 1, 1, 0, 0) == -1
    assert q(0, 64, 63, 64, 63, 12, 12, 2, 2, 11, 11) == -1   # row pitch not a multiple of 16 bytes (TMA)
    assert q(3, 64, 64, 64, 64, 12, 12, 2, 2, 11, 11) == -1   # fp64


def test_argument_errors_without_gpu(pkg):
    lib = pkg._lib.load()
    rc = lib.cno_bias_act(None, None, None, None, None, None, 0, 16, 1, 1, 0, 3, 0.2, 1.0, -1.0, None)
    assert rc > 0 and b'x and y' in lib.cno_last_error()
    rc = lib.cno_upfirdn2d(None, None, None, 0, 0, 1, 4, 4, 4, 4, 1, 1, 1, 1, 1, 1, 0, 0, 0, 1.0, None)
    assert rc > 0
    with pytest.raises(RuntimeError):
        pkg._lib.check(rc, 'upfirdn2d')


REF_SIGS = {  # reference signatures: ops/filtered_lrelu.py:48, ops/upfirdn2d.py:110,62,269,305,344, ops/bias_act.py:44
    ('filtered_lrelu', 'filtered_lrelu'): ['x', 'fu', 'fd', 'b', 'up', 'down', 'padding', 'gain', 'slope', 'clamp', 'flip_filter', 'impl'],
    ('upfirdn2d', 'upfirdn2d'): ['x', 'f', 'up', 'down', 'padding', 'flip_filter', 'gain', 'impl'],
    ('upfirdn2d', 'setup_filter'): ['f', 'device', 'normalize', 'flip_filter', 'gain', 'separable'],
    ('upfirdn2d', 'filter2d'): ['x', 'f', 'padding', 'flip_filter', 'gain', 'impl'],
    ('upfirdn2d', 'upsample2d'): ['x', 'f', 'up', 'padding', 'flip_filter', 'gain', 'impl'],
    ('upfirdn2d', 'downsample2d'): ['x', 'f', 'down', 'padding', 'flip_filter', 'gain', 'impl'],
    ('bias_act', 'bias_act'): ['x', 'b', 'dim', 'act', 'alpha', 'gain', 'clamp', 'impl'],
}


@pytest.mark.parametrize('mod,fn', list(REF_SIGS))
def test_signatures_match_reference(pkg, mod, fn):
    f = getattr(getattr(pkg, mod), fn)
    assert list(inspect.signature(f).parameters) == REF_SIGS[(mod, fn)]
    if 'impl' in REF_SIGS[(mod, fn)]:
        assert inspect.signature(f).parameters['impl'].default == 'cuda'


def test_activation_table(pkg):
    tab = pkg.bias_act.activation_funcs
    assert list(tab) == list(O.ACTIVATIONS)
    for k, (_, a, g, idx) in O.ACTIVATIONS.items():
        assert tab[k].cuda_idx == idx and abs(tab[k].def_alpha - a) < 1e-12 and abs(tab[k].def_gain - g) < 1e-12


def test_impl_ref_matches_oracle(pkg):
    g = torch.Generator().manual_seed(0)
    x = torch.randn(2, 3, 10, 12, generator=g); b = torch.randn(3, generator=g)
    cfg = O.lrelu_config(16, 8)
    x16 = torch.randn(1, 3, 16, 16, generator=g)
    a = pkg.filtered_lrelu.filtered_lrelu(x16, cfg['fu'], cfg['fd'], b, up=2, down=4, padding=cfg['padding'], impl='ref')
    r = O.filtered_lrelu(x16, cfg['fu'], cfg['fd'], b, up=2, down=4, padding=cfg['padding'])
    assert torch.allclose(a, r, atol=1e-6)
    # CPU tensors with the default impl='cuda' take the torch definition, as in the reference (filtered_lrelu.py:106-108)
    assert torch.equal(pkg.filtered_lrelu.filtered_lrelu(x16, cfg['fu'], cfg['fd'], b, up=2, down=4, padding=cfg['padding']), a)
    f = torch.randn(5, generator=g)
    assert torch.allclose(pkg.upfirdn2d.upfirdn2d(x, f, up=[2, 1], down=[1, 2], padding=[2, 1, 0, 3], impl='ref'),
                          O.upfirdn2d(x, f, up=[2, 1], down=[1, 2], padding=[2, 1, 0, 3]), atol=1e-6)
    for act in O.ACTIVATIONS:
        assert torch.allclose(pkg.bias_act.bias_act(x, b, act=act, clamp=0.8, impl='ref'), O.bias_act(x, b, act=act, clamp=0.8), atol=1e-6)
    with pytest.raises(AssertionError):
        pkg.filtered_lrelu.filtered_lrelu(x, impl='gpu')


def test_film_forward_cpu_path_equals_module(pkg, golden):
    """film_forward on a CPU tensor runs the module's own torch ops: identical to FILM.forward."""
    g = golden['cin']
    name = 'c4s16'
    C = g[f'{name}.p.norm.weight'].shape[0]

    class M(torch.nn.Module):
        pass
    m = M()
    m.norm = torch.nn.InstanceNorm2d(C, affine=True)
    m.inp2lat_sacale = torch.nn.Linear(1, 128); m.lat2scale = torch.nn.Linear(128, C)
    m.inp2lat_bias = torch.nn.Linear(1, 128); m.lat2bias = torch.nn.Linear(128, C)
    m.load_state_dict({k[len(name) + 3:]: torch.from_numpy(g[k]) for k in g.files if k.startswith(name + '.p.')})
    y = pkg.cin.film_forward(m, torch.from_numpy(g[f'{name}.x']), torch.from_numpy(g[f'{name}.t']))
    assert torch.allclose(y, torch.from_numpy(g[f'{name}.y']), atol=1e-5)


def test_film_lrelu_cpu_path_and_guards(pkg):
    """film_lrelu on CPU tensors (and for every module it has no fused kernel for) is exactly act(film(x, t)); the fused route is
    only taken for registered LReLu classes whose forward is the registered one."""
    from pregenerating_pde_b200 import cno_model
    torch.manual_seed(2)
    film = cno_model.FILM(4, dim=[2, 3], s=16)
    act = cno_model.lrelu_for(4, 16, 8)
    with torch.no_grad():
        for p_ in list(film.parameters()) + [act.bias]:
            p_.add_(0.2 * torch.randn_like(p_))
    x, t = torch.randn(2, 4, 16, 16, requires_grad=True), torch.rand(2)
    y = pkg.cin.film_lrelu(film, act, x, t)
    y2 = act(film(x, t))
    assert torch.equal(y, y2) and 'FilmLRelu' not in type(y.grad_fn).__name__
    g = torch.autograd.grad(y.square().sum(), [x] + list(film.parameters()) + [act.bias])
    g2 = torch.autograd.grad(y2.square().sum(), [x] + list(film.parameters()) + [act.bias])
    assert all(torch.equal(a, b) for a, b in zip(g, g2))
    # registration: the mirror's LReLu is registered with its own forward; a subclass or a replaced forward is not
    assert pkg.cin._lrelu_forward_of(act) is cno_model.LReLu.forward

    class Other(cno_model.LReLu):
        def forward(self, x):
            return super().forward(x) * 2
    o = Other(4, 4, 16, 8, **cno_model._filter_args(16, 8, 2.0001, 0.8))
    assert pkg.cin._lrelu_forward_of(o) is None
    assert torch.equal(pkg.cin.film_lrelu(film, o, x, t), o(film(x, t)))


def test_missing_library_is_a_hard_error(pkg, monkeypatch):
    monkeypatch.setattr(pkg._lib, '_lib', None)
    monkeypatch.setattr(pkg._lib, 'LIB_PATH', '/nonexistent/libcno_b200.so')
    with pytest.raises(pkg._lib.CnoError):
        pkg.custom_ops.get_plugin('filtered_lrelu_plugin')


def test_adam_table_record_matches_the_c_struct(pkg, tmp_path):
    """optim.MultiTensorAdamW uploads numpy records that cno_adamw_multi reads as CnoAdamTensor (include/cno_b200.h): same size, same
    field offsets as the C compiler lays the struct out"""
    import shutil, subprocess
    from pregenerating_pde_b200 import optim
    cc = shutil.which('gcc') or shutil.which('cc')
    if cc is None:
        pytest.skip('no C compiler')
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = tmp_path / 'off.c'
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "cno_b200.h"\nint main(void) { printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\\n", '
                   'sizeof(CnoAdamTensor), offsetof(CnoAdamTensor, param), offsetof(CnoAdamTensor, grad), offsetof(CnoAdamTensor, exp_avg), '
                   'offsetof(CnoAdamTensor, exp_avg_sq), offsetof(CnoAdamTensor, numel), offsetof(CnoAdamTensor, lr), '
                   'offsetof(CnoAdamTensor, weight_decay), offsetof(CnoAdamTensor, first_chunk), offsetof(CnoAdamTensor, pad)); return 0; }\n')
    exe = tmp_path / 'off'
    subprocess.check_call([cc, '-I', os.path.join(root, 'include'), str(src), '-o', str(exe)])
    got = [int(v) for v in subprocess.check_output([str(exe)]).split()]
    rec = optim._RECORD
    want = [rec.itemsize] + [rec.fields[n][1] for n in ('param', 'grad', 'exp_avg', 'exp_avg_sq', 'numel', 'lr', 'weight_decay', 'first_chunk', 'pad')]
    assert got == want


def test_optimiser_layout_helpers():
    from pregenerating_pde_b200 import optim
    a = torch.zeros(8, 4, 3, 3)
    cl = a.contiguous(memory_format=torch.channels_last)
    assert optim._dense(a) and optim._dense(cl) and not optim._dense(a[:, ::2])
    assert optim._same_layout(cl, torch.ones(8, 4, 3, 3).contiguous(memory_format=torch.channels_last))
    assert not optim._same_layout(cl, a)
    # size-1 dimensions carry arbitrary strides: they must not make equal layouts look different
    assert optim._same_layout(torch.zeros(4, 1, 3, 3), torch.zeros(4, 1, 3, 3).contiguous(memory_format=torch.channels_last))
