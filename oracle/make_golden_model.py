"""Golden vectors for the CNO_time mirror (pregenerating-pde_b200/cno_model.py), produced by the UNMODIFIED
reference model (CNO_timeModule_CIN.CNO_time, impl='ref' ops on CPU).  Build container only.

    python oracle/make_golden_model.py [/root/reference]

Stores a tiny configuration's randomised state dict, one input batch, the output, the relative-L1 loss
and the gradients of a few parameters -> tests/golden/cno_model.npz.
"""
import os, sys, warnings
import numpy as np
import torch

warnings.filterwarnings('ignore')
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg   # noqa: E402  (stub modules + REF path)

CFG = dict(in_dim=5, in_size=32, N_layers=2, N_res=1, N_res_neck=2, channel_multiplier=8, out_dim=3, is_time=1,
           nl_dim=[2, 3], latent_lift_proj_dim=16)
GRAD_KEYS = ['lift.inter_CNOBlock.activation.bias', 'encoder.0.in_norm_conditiner.norm.weight',
             'res_nets.1.in_norm_conditiner2.lat2scale.weight', 'decoder.1.convolution.weight',
             'project.convolution.bias', 'ED_expansion.0.activation.bias']
# convolution biases in front of an instance-norm CIN: mathematically zero gradient (the per-plane mean removes the bias); the
# reference's values are stored to show that they are round-off noise (tests/test_cno_model.py, cin.conv_before_film)
DEAD_BIAS_KEYS = ['encoder.0.convolution.bias', 'decoder.1.convolution.bias', 'res_nets.0.convolution1.bias',
                  'res_nets.0.convolution2.bias', 'ED_expansion.1.convolution.bias']


def main():
    mg._stub_modules()
    sys.path.insert(0, os.path.join(mg.REF, 'CNO_Experiments'))
    import CNO_timeModule_CIN as ref
    torch.manual_seed(1234)
    model = ref.CNO_time(loader_dictionary=dict(which='fpo'), **CFG)
    g = torch.Generator().manual_seed(7)
    with torch.no_grad():                    # move every parameter off its (partly zero) initial value
        for p in model.parameters():
            p.add_(0.05 * torch.randn(p.shape, generator=g))
    x = torch.randn(3, CFG['in_dim'], 32, 32, generator=g)
    t = torch.tensor([0.1, 0.45, 0.9])
    target = torch.randn(3, CFG['out_dim'], 32, 32, generator=g)
    y = model(x, t)
    loss = torch.nn.L1Loss()(target, y) / torch.nn.L1Loss()(torch.zeros_like(target), target)
    loss.backward()
    sd = model.state_dict()
    out = {'sd.' + k: v.numpy() for k, v in sd.items()}
    named = dict(model.named_parameters())
    out.update({'x': x.numpy(), 't': t.numpy(), 'target': target.numpy(), 'y': y.detach().numpy(), 'loss': np.float64(loss.item())})
    for k in GRAD_KEYS:
        out['grad.' + k] = named[k].grad.numpy()
    for k in DEAD_BIAS_KEYS:
        out['deadgrad.' + k] = named[k].grad.numpy()
    out['grad_scale'] = np.float64(max(float(p.grad.abs().max()) for p in named.values() if p.grad is not None))
    out['no_grad_params'] = np.array(sorted(k for k, p in named.items() if p.grad is None))
    out['n_params'] = np.int64(sum(p.numel() for p in model.parameters()))
    path = os.path.join(mg.OUT, 'cno_model.npz')
    if os.path.exists(path) and '--all' not in sys.argv:     # keep committed arrays (the CPU convolution backend differs by 1e-6 from run
        old = dict(np.load(path))                             # to run with the thread count); only new keys are added
        out = {**out, **old}
    np.savez_compressed(path, **out)
    print('cno_model.npz:', len(sd), 'tensors,', int(out['n_params']), 'parameters, loss', float(loss))


if __name__ == '__main__':
    main()
