"""Generate tests/golden/*.npz from the UNMODIFIED reference (run in the build container only).

    python oracle/make_golden.py [/root/reference]

Imports /root/reference/CNO_Experiments as-is (with stub modules for the third-party imports the
hot path does not use: pytorch_lightning, matplotlib, h5py, netCDF4), runs the reference's own
impl='ref' code on seeded inputs and stores inputs + outputs.  The GPU box has no /root/reference;
tests only read the committed .npz files.
"""
import os, sys, types, math, warnings
import numpy as np
import torch

warnings.filterwarnings('ignore')
REF = sys.argv[1] if len(sys.argv) > 1 else '/root/reference'
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', 'tests', 'golden')


def _stub_modules():
    pl = types.ModuleType('pytorch_lightning')
    class LightningModule(torch.nn.Module):
        def log(self, *a, **k): pass
        def save_hyperparameters(self, *a, **k): pass
    pl.LightningModule = LightningModule
    sys.modules['pytorch_lightning'] = pl
    for name in ['matplotlib', 'matplotlib.pyplot', 'h5py', 'netCDF4']:
        sys.modules[name] = types.ModuleType(name)
    sys.modules['matplotlib'].pyplot = sys.modules['matplotlib.pyplot']


def main():
    _stub_modules()
    sys.path.insert(0, os.path.join(REF, 'CNO_Experiments'))
    from torch_utils.ops import filtered_lrelu as r_fl, upfirdn2d as r_up, bias_act as r_ba
    from training.filtered_networks import LReLu
    import CNO_timeModule_CIN as r_cin
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(os.cpu_count())

    # ---- filtered_lrelu: the three CNO configs through the reference LReLu module ----------
    fl = {}
    cases = [('I16', 16, 16, 3, 2), ('D16', 16, 8, 3, 2), ('U8', 8, 16, 3, 2), ('I32', 32, 32, 2, 1),
             ('D32', 32, 16, 2, 1), ('U16', 16, 32, 2, 1), ('I8', 8, 8, 4, 2)]
    for name, s_in, s_out, C, N in cases:
        m = LReLu(C, C, s_in, s_out, s_in, s_out, s_in / 2.0001, s_out / 2.0001,
                  0.8 * s_in - s_in / 2.0001, 0.8 * s_out - s_out / 2.0001)
        g = torch.Generator().manual_seed(hash(name) % 1000 if False else len(name) * 31 + s_in)
        x = torch.randn(N, C, s_in, s_in, generator=g).requires_grad_(True)
        b = torch.randn(C, generator=g).requires_grad_(True)
        dy = torch.randn(N, C, s_out, s_out, generator=g)
        kw = dict(fu=m.up_filter, fd=m.down_filter, up=m.up_factor, down=m.down_factor, padding=m.padding,
                  gain=math.sqrt(2), slope=0.2, clamp=None)
        y = r_fl._filtered_lrelu_ref(x, b=b, **kw)
        dx, db = torch.autograd.grad(y, [x, b], dy)
        v = r_up._upfirdn2d_ref(r_ba._bias_act_ref(x.detach(), b.detach()), m.up_filter, up=m.up_factor,
                                padding=m.padding, gain=m.up_factor ** 2) * math.sqrt(2)
        for k, t in dict(x=x, b=b, dy=dy, y=y, dx=dx, db=db, v=v, fu=m.up_filter, fd=m.down_filter).items():
            fl[f'{name}.{k}'] = t.detach().numpy()
        fl[f'{name}.cfg'] = np.array([m.up_factor, m.down_factor] + list(m.padding), dtype=np.int64)
    # generic arguments: clamp, flip_filter, asymmetric filters, non-square padding
    g = torch.Generator().manual_seed(7)
    x = torch.randn(2, 3, 12, 10, generator=g).requires_grad_(True)
    b = torch.randn(3, generator=g).requires_grad_(True)
    fu = torch.randn(8, generator=g); fd = torch.randn(8, generator=g)
    kw = dict(fu=fu, fd=fd, up=2, down=2, padding=[5, 4, 6, 3], gain=1.3, slope=0.1, clamp=0.7, flip_filter=True)
    y = r_fl._filtered_lrelu_ref(x, b=b, **kw)
    dy = torch.randn(y.shape, generator=g)
    dx, db = torch.autograd.grad(y, [x, b], dy)
    for k, t in dict(x=x, b=b, dy=dy, y=y, dx=dx, db=db, fu=fu, fd=fd).items():
        fl[f'GEN.{k}'] = t.detach().numpy()
    np.savez_compressed(os.path.join(OUT, 'filtered_lrelu.npz'), **fl)

    # ---- upfirdn2d ---------------------------------------------------------------------------
    up = {}
    g = torch.Generator().manual_seed(11)
    specs = [('up2', dict(up=2, down=1, padding=[6, 5, 6, 5], gain=4), 12),
             ('down2', dict(up=1, down=2, padding=[5, 5, 5, 5], gain=1), 12),
             ('xy', dict(up=[2, 1], down=[1, 3], padding=[3, 2, 4, 1], gain=1.5, flip_filter=True), 7),
             ('f2d', dict(up=2, down=2, padding=[2, 1, 0, 3], gain=1), (3, 5)),
             ('crop', dict(up=1, down=1, padding=[-1, -2, 1, -1], gain=1), 3)]
    for name, kw, taps in specs:
        x = torch.randn(2, 3, 9, 11, generator=g).requires_grad_(True)
        f = torch.randn(taps, generator=g) if isinstance(taps, int) else torch.randn(*taps, generator=g)
        y = r_up._upfirdn2d_ref(x, f, **kw)
        dy = torch.randn(y.shape, generator=g)
        dx, = torch.autograd.grad(y, [x], dy)
        for k, t in dict(x=x, f=f, y=y, dy=dy, dx=dx).items():
            up[f'{name}.{k}'] = t.detach().numpy()
    np.savez_compressed(os.path.join(OUT, 'upfirdn2d.npz'), **up)

    # ---- bias_act: all nine activations ------------------------------------------------------
    ba = {}
    g = torch.Generator().manual_seed(13)
    x = torch.randn(2, 5, 4, 6, generator=g)
    b = torch.randn(5, generator=g)
    dy = torch.randn(2, 5, 4, 6, generator=g)
    ba['x'], ba['b'], ba['dy'] = x.numpy(), b.numpy(), dy.numpy()
    for act in r_ba.activation_funcs:
        for tag, kw in [('def', {}), ('arg', dict(alpha=0.3, gain=1.7, clamp=0.9))]:
            xx = x.clone().requires_grad_(True); bb = b.clone().requires_grad_(True)
            y = r_ba._bias_act_ref(xx, bb, act=act, **kw)
            dx, db = torch.autograd.grad(y, [xx, bb], dy)
            ba[f'{act}.{tag}.y'], ba[f'{act}.{tag}.dx'], ba[f'{act}.{tag}.db'] = y.detach().numpy(), dx.numpy(), db.numpy()
    np.savez_compressed(os.path.join(OUT, 'bias_act.npz'), **ba)

    # ---- FILM / CIN --------------------------------------------------------------------------
    ci = {}
    torch.manual_seed(17)
    for name, C, S, N in [('c4s16', 4, 16, 3), ('c6s8', 6, 8, 2)]:
        film = r_cin.FILM(C, dim=[2, 3], s=S)
        with torch.no_grad():
            for p in film.parameters():
                p.copy_(torch.randn_like(p) * 0.5)
        x = (torch.randn(N, C, S, S) * 2 + 0.5).requires_grad_(True)
        t = torch.rand(N)
        y = film(x, t)
        dy = torch.randn_like(y)
        params = list(film.parameters())
        grads = torch.autograd.grad(y, [x] + params, dy)
        ci[f'{name}.x'], ci[f'{name}.t'], ci[f'{name}.y'], ci[f'{name}.dy'], ci[f'{name}.dx'] = \
            x.detach().numpy(), t.numpy(), y.detach().numpy(), dy.numpy(), grads[0].numpy()
        for (pn, p), gr in zip(film.named_parameters(), grads[1:]):
            ci[f'{name}.p.{pn}'] = p.detach().numpy()
            ci[f'{name}.g.{pn}'] = gr.numpy()
    np.savez_compressed(os.path.join(OUT, 'cin.npz'), **ci)
    for fn in sorted(os.listdir(OUT)):
        print(fn, os.path.getsize(os.path.join(OUT, fn)))


if __name__ == '__main__':
    main()
