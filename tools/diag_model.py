"""GPU: CNO_time mirror vs the golden reference gradients in three modes, to separate kernel error from conv-backend round-off."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import pregenerating_pde_b200 as pkg
from pregenerating_pde_b200 import cno_model, cin
import test_cno_model as T
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
g = np.load(os.path.join(ROOT, 'tests/golden/cno_model.npz'))
sd = {k[3:]: torch.from_numpy(g[k]) for k in g.files if k.startswith('sd.')}

def eager_film(self, x, t, residual=None):
    t = t.reshape(-1, 1).to(x.dtype)
    s = self.lat2scale(self.inp2lat_sacale(t)); q = self.lat2bias(self.inp2lat_bias(t))
    y = self.norm(x) * s[:, :, None, None] + q[:, :, None, None]
    return y if residual is None else residual + y

def run(mode, dtype=torch.float32):
    m = cno_model.CNO_time(**T.CFG).to('cuda'); m.load_state_dict(sd)
    if mode == 'ref':
        for mod in m.modules():
            if isinstance(mod, cno_model.LReLu): mod.impl = 'ref'
        cno_model.FILM.forward = eager_film
    else:
        cno_model.FILM.forward = cin.film_forward
        cin.FUSE_LINEARS = (mode == 'fused')
    m = m.to(dtype)
    x, t, target = (torch.from_numpy(g[k]).to('cuda').to(dtype) for k in ('x', 't', 'target'))
    y = m(x, t); loss = m.training_loss(y, target); loss.backward()
    named = dict(m.named_parameters())
    errs = sorted(((T._rel(named[k[5:]].grad, g[k]), k) for k in g.files if k.startswith('grad.')), reverse=True)
    print(mode, dtype, 'y', T._rel(y, g['y']), 'top:', [(f'{e:.1e}', k[5:]) for e, k in errs[:6]])

for mode in ('fused', 'cin', 'ref'):
    run(mode)

