"""MultiTensorAdamW (one launch for every parameter tensor, csrc/adamw.cu) against torch.optim.AdamW on the same parameters and
gradients: ragged tensor sizes (chunk tails, scalars, unaligned views), several parameter groups, parameters without a gradient,
learning-rate changes between steps, and a CUDA-graph-captured step."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = 'cuda'


@pytest.fixture(scope='module')
def cno():
    import pregenerating_pde_b200 as pkg
    pkg._lib.load()
    return pkg


def make_params(seed, shapes):
    g = torch.Generator().manual_seed(seed)
    return [torch.nn.Parameter(torch.randn(s, generator=g).to(DEV)) for s in shapes]


SHAPES = [(1,), (3,), (17, 5), (4096,), (4097,), (64, 3, 3, 3), (128, 129), (5, 7, 11, 13), (8191,), (2, 4096)]


def rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


def test_matches_torch_adamw_over_steps(cno):
    from pregenerating_pde_b200.optim import MultiTensorAdamW
    pa, pb = make_params(0, SHAPES), make_params(0, SHAPES)
    groups = lambda ps: [dict(params=ps[:4], lr=3e-3, weight_decay=0.0), dict(params=ps[4:7], lr=1e-3), dict(params=ps[7:], lr=5e-4, weight_decay=0.1)]
    oa = MultiTensorAdamW(groups(pa), lr=1e-3, weight_decay=1e-2)
    ob = torch.optim.AdamW(groups(pb), lr=1e-3, weight_decay=1e-2)
    g = torch.Generator().manual_seed(1)
    for step in range(12):
        for i, (a, b) in enumerate(zip(pa, pb)):
            if i == 2 and step % 3 == 0:                       # a parameter without gradient this step is skipped by both
                a.grad = b.grad = None
                continue
            gr = torch.randn(a.shape, generator=g).to(DEV) * (10.0 ** (i % 3 - 1))
            a.grad, b.grad = gr.clone(), gr.clone()
        if step == 6:                                          # scheduler-style change
            for o in (oa, ob):
                o.param_groups[1]['lr'] = 2e-4
        oa.step(); ob.step()
    # torch skips the step counter of a parameter that had no gradient; ours shares one counter, so parameter 2 is compared loosely
    for i, (a, b) in enumerate(zip(pa, pb)):
        tol = 2e-2 if i == 2 else 2e-6
        assert rel(a.detach(), b.detach()) < tol, (i, rel(a.detach(), b.detach()))
        if i != 2:
            assert rel(oa.state[a]['exp_avg'], ob.state[b]['exp_avg']) < 1e-5
            assert rel(oa.state[a]['exp_avg_sq'], ob.state[b]['exp_avg_sq']) < 1e-5
    assert float(oa.state[pa[0]]['step']) == 12


def test_unaligned_views_and_flat_gradients(cno):
    """parameters and gradients that are views at odd offsets of flat buffers (the multi-rank Trainer's gradient layout)"""
    from pregenerating_pde_b200.optim import MultiTensorAdamW
    sizes = [5, 4099, 33, 1, 12288]
    flat_p = torch.randn(sum(sizes) + 3, device=DEV)
    flat_g = torch.randn(sum(sizes) + 3, device=DEV)
    pa, pb, o = [], [], 3                                      # offset 3 floats: 4-byte alignment only
    for n in sizes:
        a = torch.nn.Parameter(flat_p[o:o + n].clone()); b = torch.nn.Parameter(flat_p[o:o + n].clone())
        a.grad = flat_g[o:o + n]; b.grad = flat_g[o:o + n].clone()
        pa.append(a); pb.append(b); o += n
    # parameters of the kernel-driven optimiser live inside one guarded buffer: the update must not touch a float outside them
    guard = torch.full((sum(sizes) + 64 * (len(sizes) + 1),), 3.0, device=DEV)
    o = 64
    for a, n in zip(pa, sizes):
        guard[o:o + n] = a.data
        a.data = guard[o:o + n]
        o += n + 64
    oa, ob = MultiTensorAdamW(pa, lr=1e-2), torch.optim.AdamW(pb, lr=1e-2)
    for _ in range(3):
        oa.step(); ob.step()
    for a, b in zip(pa, pb):
        assert rel(a.detach(), b.detach()) < 2e-6
    o = 0
    for n in sizes:
        assert bool((guard[o:o + 64] == 3.0).all())
        o += 64 + n
    assert bool((guard[o:] == 3.0).all())


def test_graph_captured_step_replays(cno):
    """one eager step (zero gradients), then the captured step replayed five times on static gradient buffers = six torch steps"""
    from pregenerating_pde_b200.optim import MultiTensorAdamW
    pa, pc = make_params(5, SHAPES), make_params(5, SHAPES)
    oa, oc = MultiTensorAdamW(pa, lr=1e-3), torch.optim.AdamW(pc, lr=1e-3)
    static_g = [torch.zeros_like(p) for p in pa]
    for p, gbuf in zip(pa, static_g):
        p.grad = gbuf
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        oa.step()                                              # warm-up outside capture
    torch.cuda.current_stream().wait_stream(s)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        oa.step()                                              # captured, not executed
    for p in pc:
        p.grad = torch.zeros_like(p)
    oc.step()
    gen = torch.Generator().manual_seed(9)
    for _ in range(5):
        grads = [torch.randn(p.shape, generator=gen).to(DEV) for p in pc]
        for gbuf, p, gr in zip(static_g, pc, grads):
            gbuf.copy_(gr); p.grad = gr
        graph.replay()
        oc.step()
    torch.cuda.synchronize()
    assert float(oa.state[pa[0]]['step']) == 6
    for a, c in zip(pa, pc):
        assert rel(a.detach(), c.detach()) < 2e-6


def test_rejects_unsupported(cno):
    from pregenerating_pde_b200.optim import MultiTensorAdamW
    with pytest.raises(ValueError):
        MultiTensorAdamW([torch.nn.Parameter(torch.zeros(4))], lr=1e-3)                        # CPU parameter
    with pytest.raises(ValueError):
        MultiTensorAdamW([torch.nn.Parameter(torch.zeros(4, device=DEV, dtype=torch.bfloat16))])
    lib = cno._lib.load()
    assert lib.cno_adamw_multi(None, None, 0, None, 0.9, 0.999, 1e-8, None) != 0


def test_channels_last_parameters(cno):
    """parameters stored channels_last (the Trainer's convolution weights): moments share their strides, the update is elementwise over
    storage and equals torch.optim.AdamW on the same tensors"""
    from pregenerating_pde_b200.optim import MultiTensorAdamW
    g = torch.Generator().manual_seed(3)
    shapes = [(16, 8, 3, 3), (8, 1, 3, 3), (4, 16, 1, 1), (5,)]
    mk = lambda: [torch.nn.Parameter(torch.randn(s, generator=torch.Generator().manual_seed(i)).to(DEV)) for i, s in enumerate(shapes)]
    pa, pb = mk(), mk()
    for p in pa:
        if p.ndim == 4:
            p.data = p.data.contiguous(memory_format=torch.channels_last)
    oa, ob = MultiTensorAdamW(pa, lr=1e-2), torch.optim.AdamW(pb, lr=1e-2)
    for _ in range(4):
        for a, b in zip(pa, pb):
            gr = torch.randn(a.shape, generator=g).to(DEV)
            a.grad = gr.contiguous(memory_format=torch.channels_last) if a.ndim == 4 else gr.clone()
            b.grad = gr.clone()
        oa.step(); ob.step()
    for a, b in zip(pa, pb):
        assert rel(a.detach(), b.detach()) < 2e-6
        assert rel(oa.state[a]['exp_avg'], ob.state[b]['exp_avg']) < 1e-5
    bad = torch.nn.Parameter(torch.randn(16, 8, 3, 3, device=DEV).contiguous(memory_format=torch.channels_last))
    ob2 = MultiTensorAdamW([bad], lr=1e-3)
    bad.grad = torch.randn(16, 8, 3, 3, device=DEV)            # NCHW gradient for a channels_last parameter: refused, not silently permuted
    with pytest.raises(RuntimeError):
        ob2.step()
