"""Fine-tune wrappers (finetune.py) against the UNMODIFIED reference (test_and_fine_tune_utils/fine_tune_lift.py,
CNO_timeModule_CIN.py:983-994) on CPU: identical state_dict keys, identical forward on a shared state dict, identical
optimiser grouping.  Build container only (the reference does not travel to the GPU box)."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = '/root/reference/CNO_Experiments'

_SCRIPT = r'''
import sys, types, torch
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
from test_and_fine_tune_utils import fine_tune_lift as RF
import pregenerating_pde_b200 as cno
from pregenerating_pde_b200 import cno_model, finetune

for (new_in, new_out) in [(4, 3), (5, 4), (4, 4)]:
    torch.manual_seed(0)
    kw = dict(in_dim=5, in_size=128, N_layers=1, N_res=1, N_res_neck=1, channel_multiplier=4, out_dim=4, is_time=1, nl_dim=[2, 3])
    ref = R.CNO_time(batch_norm=False, **kw)
    ref = RF.initialize_FT(ref, old_in_dim=5, new_in_dim=new_in, new_out_dim=new_out, old_out_dim=4).cpu()
    ours = finetune.initialize_FT(cno_model.CNO_time(impl='ref', **kw), 5, new_in, new_out, 4)
    sd = ref.state_dict()
    assert set(sd.keys()) == set(ours.state_dict().keys()), set(sd.keys()) ^ set(ours.state_dict().keys())
    with torch.no_grad():
        for k, v in sd.items():
            if v.dtype.is_floating_point and 'filter' not in k:
                v.add_(0.05 * torch.randn_like(v))
    ours.load_state_dict(sd)
    x = torch.randn(2, new_in, 128, 128); t = torch.rand(2)
    yr, yo = ref(x, t), ours(x, t)
    assert yr.shape == yo.shape, (yr.shape, yo.shape)
    err = float((yr - yo).abs().max() / yr.abs().max())
    assert err < 2e-5, err
    # the three AdamW groups (CNO_timeModule_CIN.py:983-994)
    names = dict(ours.named_parameters())
    inv = {id(p): n for n, p in names.items()}
    groups = finetune.finetune_param_groups(ours, 1e-4, 1e-3, 1e-2)
    got = [sorted(inv[id(p)] for p in g['params']) for g in groups]
    rn = [n for n, _ in ref.named_parameters()]
    exp = [sorted(n for n in rn if ('project' not in n) and ('lift' not in n) and ('in_norm_conditiner' not in n)),
           sorted(n for n in rn if (('project' in n) or ('lift' in n)) and ('in_norm_conditiner' not in n)),
           sorted(n for n in rn if ('in_norm_conditiner' in n) and ('project' not in n) and ('lift' not in n))]
    assert got == exp
    assert [g['lr'] for g in groups] == [1e-4, 1e-3, 1e-2]
print('FT_OK')
'''


@pytest.mark.skipif(not os.path.isdir(REF), reason='reference checkout not present (GPU box)')
def test_finetune_wrappers_match_reference():
    script = f'ROOT = {ROOT!r}\nREF = {REF!r}\n' + _SCRIPT
    r = subprocess.run([sys.executable, '-c', script], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and 'FT_OK' in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_wide_finetune_config_shape():
    """BASELINE.json configs[4]: features up to 512 at 256^2, ~67.7 M parameters, 128^2 output through the fresh projection."""
    import pregenerating_pde_b200 as cno
    from pregenerating_pde_b200 import cno_model, finetune
    ft = finetune.wide_finetune_config()
    with torch.device('meta'):
        m = finetune.initialize_FT(cno_model.CNO_time(impl='ref', **ft['model']), **ft['ft'])
    assert m.encoder_features == [32, 64, 128, 256, 512] and m.encoder_sizes == [256, 128, 64, 32, 16]
    n = sum(p.numel() for p in m.parameters())
    assert 67_000_000 < n < 68_500_000, n
    assert int(m.project.project.inter_CNOBlock.activation.out_size[0]) == 128
