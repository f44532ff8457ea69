"""Golden vectors for the radial (jinc) down-filter flavour of LReLu (`use_radial_filters=True`,
training/filtered_networks.py:245-247, :293-301), produced by the UNMODIFIED reference LReLu with impl='ref' on CPU.
Build container only.   python oracle/make_golden_radial.py [/root/reference]   -> tests/golden/lrelu_radial.npz
Test infrastructure: only tests/ read the result."""
import os, sys, warnings
import numpy as np
import torch

warnings.filterwarnings('ignore')
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg   # noqa: E402

CASES = {'I': (16, 16), 'D': (16, 8), 'U': (8, 16)}


def main():
    mg._stub_modules()
    sys.path.insert(0, os.path.join(mg.REF, 'CNO_Experiments'))
    from training.filtered_networks import LReLu
    out = {}
    g = torch.Generator().manual_seed(3)
    for name, (si, so) in CASES.items():
        kw = dict(in_sampling_rate=si, out_sampling_rate=so, in_cutoff=si / 2.0001, out_cutoff=so / 2.0001,
                  in_half_width=0.8 * si - si / 2.0001, out_half_width=0.8 * so - so / 2.0001)
        m = LReLu(in_channels=3, out_channels=3, in_size=si, out_size=so, use_radial_filters=True, **kw)   # CPU tensors -> the reference's impl='ref' definition
        with torch.no_grad():
            m.bias.copy_(torch.randn(3, generator=g))
        x = torch.randn(2, 3, si, si, generator=g, requires_grad=True)
        y = m(x)
        dy = torch.randn(y.shape, generator=g)
        dx, db = torch.autograd.grad(y, [x, m.bias], dy)
        out.update({f'{name}.x': x.detach().numpy(), f'{name}.b': m.bias.detach().numpy(), f'{name}.dy': dy.numpy(),
                    f'{name}.y': y.detach().numpy(), f'{name}.dx': dx.numpy(), f'{name}.db': db.numpy(),
                    f'{name}.fu': m.up_filter.numpy(), f'{name}.fd': m.down_filter.numpy(),
                    f'{name}.padding': np.array(m.padding)})
    np.savez_compressed(os.path.join(mg.OUT, 'lrelu_radial.npz'), **out)
    print('lrelu_radial.npz:', {k: v.shape for k, v in out.items() if k.endswith('.fd')})


if __name__ == '__main__':
    main()
