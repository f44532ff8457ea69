"""The CNO_time mirror (callers of the hot path) against golden vectors produced by the unmodified reference
model (oracle/make_golden_model.py): identical state_dict keys, forward, loss and gradients.
CPU: the mirror runs impl='ref'-equivalent torch ops (no CUDA tensor -> the package's ref definitions);
GPU: the same checkpoint through the sm_100a kernels."""
import os, sys
import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
CFG = dict(in_dim=5, in_size=32, N_layers=2, N_res=1, N_res_neck=2, channel_multiplier=8, out_dim=3, is_time=1,
           nl_dim=[2, 3], latent_lift_proj_dim=16)


def _load(golden):
    g = golden['cno_model']
    sd = {k[3:]: torch.from_numpy(g[k]) for k in g.files if k.startswith('sd.')}
    return g, sd


def _build(dev):
    import pregenerating_pde_b200 as pkg
    from pregenerating_pde_b200 import cno_model
    return cno_model.CNO_time(**CFG).to(dev)


def _rel(a, b):
    a, b = torch.as_tensor(a).detach().double().cpu(), torch.as_tensor(b).detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def _eager_film(self, x, t, residual=None):
    """FILM.forward in plain torch ops (CNO_timeModule_CIN.py:60-70); residual: the mirror's ResidualBlock epilogue"""
    t = t.reshape(-1, 1).to(x.dtype)
    s, q = self.lat2scale(self.inp2lat_sacale(t)), self.lat2bias(self.inp2lat_bias(t))
    y = self.norm(x) * s[:, :, None, None] + q[:, :, None, None]
    return y if residual is None else residual + y


def _torch_backend_deviation(golden, dev):
    """gradients of the same checkpoint on `dev` with torch ops only (impl='ref' LReLu, eager FILM): their distance to the CPU
    golden gradients is what the device's convolution backend alone contributes (none of our kernels runs)."""
    from pregenerating_pde_b200 import cno_model
    g, sd = _load(golden)
    model = _build(dev)
    model.load_state_dict(sd, strict=True)
    for m in model.modules():
        if isinstance(m, cno_model.LReLu):
            m.impl = 'ref'
    fused = cno_model.FILM.forward
    cno_model.FILM.forward = _eager_film
    try:
        x, t, target = (torch.from_numpy(g[k]).to(dev) for k in ('x', 't', 'target'))
        model.training_loss(model(x, t), target).backward()
    finally:
        cno_model.FILM.forward = fused
    named = dict(model.named_parameters())
    return {k: _rel(named[k[5:]].grad, g[k]) for k in g.files if k.startswith('grad.')}


def _check(golden, dev, tol, backend_dev=None):
    g, sd = _load(golden)
    model = _build(dev)
    assert sorted(model.state_dict().keys()) == sorted(sd.keys())            # checkpoint compatibility
    model.load_state_dict(sd, strict=True)
    assert sum(p.numel() for p in model.parameters()) == int(g['n_params'])
    x, t, target = (torch.from_numpy(g[k]).to(dev) for k in ('x', 't', 'target'))
    y = model(x, t)
    assert _rel(y, g['y']) < tol
    loss = model.training_loss(y, target)
    assert abs(float(loss.detach()) - float(g['loss'])) / float(g['loss']) < tol
    loss.backward()
    named = dict(model.named_parameters())
    for k in g.files:
        if k.startswith('grad.'):
            # LReLu bias gradients are sums of dx over N*H*W with heavy cancellation (|db| << sum|dx|), so the fp32 round-off of
            # a different convolution backend upstream shows up amplified there; db itself is pinned at op level to 1e-5
            lim = 100 * tol if (k.endswith('activation.bias') and dev != 'cpu') else 10 * tol
            if backend_dev is not None:
                # measured on B200: with torch ops ONLY the GPU already sits 2.3e-3 (lift bias) / 7.8e-4 (first norm weight) from
                # the CPU golden on two cancellation-heavy sums (the L1 loss gradient is sign(pred - target): one flipped sign
                # moves them), identically for the fused and the unfused path; allow what the backend alone produces
                lim = max(lim, 2 * backend_dev[k])
            assert _rel(named[k[5:]].grad, g[k]) < lim, (k, _rel(named[k[5:]].grad, g[k]))
    # the reference builds decoder_inv[N_layers] but never calls it: same parameters stay without gradient
    assert sorted(k for k, p in named.items() if p.grad is None) == list(g['no_grad_params'])


@pytest.fixture
def keep_conv_bias():
    """the golden comparison runs the convolutions WITH their bias, like the reference (bit-identical convolution path on the CPU);
    the bias-free form has its own test below"""
    from pregenerating_pde_b200 import cin
    prev, cin.SKIP_DEAD_BIAS = cin.SKIP_DEAD_BIAS, False
    yield
    cin.SKIP_DEAD_BIAS = prev


def test_mirror_matches_reference_cpu(golden, keep_conv_bias):
    _check(golden, 'cpu', 2e-5)


def _dead_bias_check(golden, dev, tol):
    """cin.conv_before_film: a convolution bias in front of an instance-norm CIN vanishes in the per-plane mean, so the package
    runs the convolution without it.  Claims checked: (1) the reference's OWN gradient of such a bias (golden file) is round-off
    noise, (2) output and all other gradients are unchanged.  (2) uses a smooth loss: the model's relative-L1 loss has the
    gradient sign(pred - target), where a 1e-7 change of pred flips signs and moves cancellation-heavy sums by 1e-3."""
    from pregenerating_pde_b200 import cin
    g, sd = _load(golden)
    gscale = float(g['grad_scale'])
    x, t, target = (torch.from_numpy(g[k]).to(dev) for k in ('x', 't', 'target'))
    res = {}
    prev = cin.SKIP_DEAD_BIAS
    try:
        for skip in (False, True):
            cin.SKIP_DEAD_BIAS = skip
            model = _build(dev)
            model.load_state_dict(sd, strict=True)
            y = model(x, t)
            ((y - target) ** 2).mean().backward()
            res[skip] = (y.detach(), {k: p.grad for k, p in model.named_parameters()})
    finally:
        cin.SKIP_DEAD_BIAS = prev
    assert _rel(res[True][0], res[False][0].cpu().numpy()) < tol
    gs = max(float(v.abs().max()) for v in res[False][1].values() if v is not None)
    dead = [k for k, v in res[True][1].items() if v is None and res[False][1][k] is not None]
    assert len(dead) >= 10 and all(k.endswith(('convolution.bias', 'convolution1.bias', 'convolution2.bias')) for k in dead)
    stored = [k[9:] for k in g.files if k.startswith('deadgrad.')]
    assert len(stored) >= 5 and set(stored) <= set(dead)
    for k in stored:                                           # the unmodified reference's own gradients of these biases
        assert float(abs(g['deadgrad.' + k]).max()) < 1e-5 * gscale, (k, 'reference gradient is not noise')
    for k in dead:
        assert float(res[False][1][k].abs().max()) < 1e-5 * gs
    # all other gradients: equal up to the leaky-ReLU kinks (a pre-activation within 1e-7 of zero changes slope 1 <-> 0.2 when the
    # rounding of the convolution changes; each such flip is a finite jump of the gradient), hence an L2 and a loose max bound
    num = sum(float(((v - res[False][1][k]) ** 2).sum()) for k, v in res[True][1].items() if v is not None)
    den = sum(float((res[False][1][k] ** 2).sum()) for k, v in res[True][1].items() if v is not None)
    assert (num / den) ** 0.5 < 2e-3, (num / den) ** 0.5
    for k, v in res[True][1].items():
        if v is not None:
            assert float((v - res[False][1][k]).abs().max()) < 1e-2 * gs, k


def test_dead_conv_bias_is_exact_algebra_cpu(golden):
    _dead_bias_check(golden, 'cpu', 2e-5)


@pytest.mark.gpu
def test_dead_conv_bias_is_exact_algebra_gpu(golden):
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    _dead_bias_check(golden, 'cuda', 5e-5)


@pytest.mark.gpu
def test_mirror_matches_reference_gpu(golden, keep_conv_bias):
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    _check(golden, 'cuda', 5e-5, backend_dev=_torch_backend_deviation(golden, 'cuda'))


@pytest.mark.skipif(not os.path.isdir('/root/reference/CNO_Experiments'), reason='reference checkout not present')
def test_mirror_key_set_full_config():
    """full TrainCNO_time_L configuration: same parameter names / shapes as the reference (18,280,147 parameters)."""
    import make_golden as mg
    mg._stub_modules()
    sys.path.insert(0, '/root/reference/CNO_Experiments')
    import CNO_timeModule_CIN as ref
    from pregenerating_pde_b200 import cno_model
    cfg = cno_model.train_config_L()
    a = ref.CNO_time(loader_dictionary={}, **cfg).state_dict()
    b = cno_model.CNO_time(**cfg).state_dict()
    assert {k: tuple(v.shape) for k, v in a.items()} == {k: tuple(v.shape) for k, v in b.items()}
    assert sum(v.numel() for k, v in cno_model.CNO_time(**cfg).named_parameters()) == 18280147


# ---- radial (jinc) down filter flavour of LReLu: golden vectors from the unmodified reference LReLu -------------
def _radial_case(golden, name, dev):
    from pregenerating_pde_b200 import cno_model
    g = golden['lrelu_radial']
    si, so = {'I': (16, 16), 'D': (16, 8), 'U': (8, 16)}[name]
    m = cno_model.lrelu_for(3, si, so, use_radial_filters=True).to(dev)
    assert m.down_filter.ndim == 2 and m.padding == [int(v) for v in g[f'{name}.padding']]
    assert _rel(m.up_filter, g[f'{name}.fu']) < 1e-6 and _rel(m.down_filter, g[f'{name}.fd']) < 1e-6
    with torch.no_grad():
        m.bias.copy_(torch.from_numpy(g[f'{name}.b']))
    x = torch.from_numpy(g[f'{name}.x']).to(dev).requires_grad_(True)
    y = m(x)
    dx, db = torch.autograd.grad(y, [x, m.bias], torch.from_numpy(g[f'{name}.dy']).to(dev))
    assert _rel(y, g[f'{name}.y']) < 1e-5 and _rel(dx, g[f'{name}.dx']) < 1e-5 and _rel(db, g[f'{name}.db']) < 1e-5


@pytest.mark.parametrize('name', ['I', 'D', 'U'])
def test_lrelu_radial_filters_cpu(golden, name):
    _radial_case(golden, name, 'cpu')


@pytest.mark.gpu
@pytest.mark.parametrize('name', ['I', 'D', 'U'])
def test_lrelu_radial_filters_gpu(golden, name):
    """2-D down filter: no fused variant -> generic composition (fast separable up pass, act kernel, generic 2-D down pass)"""
    _radial_case(golden, name, 'cuda')
