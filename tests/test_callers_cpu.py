"""CPU tests of the callers either side of the hot path that SURVEY 8f-4 lists: the torch-only `LReLu_torch` activation
(`activation='cno_lrelu_torch'`) and the device-resident sample construction of NSFlowTimeDataset, each against the unmodified
reference when /root/reference is present (build container) and against a plain restatement otherwise."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = '/root/reference/CNO_Experiments'

_SCRIPT = r'''
import sys, types, torch, numpy as np
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
from training.filtered_networks import LReLu_torch as RefAct
import CNO_timeModule_CIN as R
import pregenerating_pde_b200 as cno
from pregenerating_pde_b200 import cno_model

torch.manual_seed(0)
for (i, o) in [(16, 16), (16, 8), (8, 16)]:
    a, b = RefAct(5, 5, i, o, i, o), cno_model.LReLu_torch(5, 5, i, o, i, o)
    assert set(a.state_dict().keys()) == set(b.state_dict().keys())
    with torch.no_grad():
        a.bias.normal_()
    b.load_state_dict(a.state_dict())
    x = torch.randn(2, 5, i, i)
    assert torch.allclose(a(x), b(x), atol=1e-6), (i, o)
# a block that selects it: same keys, same forward
kw = dict(in_channels=3, out_channels=4, in_size=16, out_size=8, is_time=True, nl_dim=[2, 3], activation='cno_lrelu_torch')
ra = R.CNOBlock(batch_norm=False, **kw)
ob = cno_model.CNOBlock(3, 4, 16, 8, is_time=True, nl_dim=[2, 3], activation='cno_lrelu_torch')
assert set(ra.state_dict().keys()) == set(ob.state_dict().keys())
ob.load_state_dict(ra.state_dict())
x, t = torch.randn(2, 3, 16, 16), torch.rand(2)
assert torch.allclose(ra(x, t), ob(x, t), atol=1e-5)
print('OK')
'''


@pytest.mark.skipif(not os.path.isdir(REF), reason='the reference checkout exists in the build container only')
def test_lrelu_torch_matches_reference():
    code = f'ROOT = {ROOT!r}\nREF = {REF!r}\n' + _SCRIPT
    r = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and 'OK' in r.stdout, r.stdout + r.stderr


def test_lrelu_torch_shapes_and_grad():
    import pregenerating_pde_b200 as cno
    act = cno.cno_model.LReLu_torch(3, 3, 12, 6)
    x = torch.randn(2, 3, 12, 12, requires_grad=True)
    y = act(x)
    assert y.shape == (2, 3, 6, 6)
    y.sum().backward()
    assert x.grad is not None and act.bias.grad.shape == (3,)
    with pytest.raises(ValueError):
        cno.cno_model.CNOBlock(3, 3, 8, 8, activation='lrelu')          # LReLu_standard is not part of the mirrored path


def test_device_batcher_matches_getitem_restatement():
    """NSFlowTimeDataset.__getitem__ (CNO_TimeLoaders.py:219-248) restated per sample on the host vs the batched construction."""
    import pregenerating_pde_b200 as cno
    rng = np.random.default_rng(0)
    N, T, S = 6, 21, 8
    data = rng.standard_normal((N, T, S, S, 6)).astype(np.float32)
    time_indices = [(0, 2 * j) for j in range(0, 8)]                        # allowed_transitions None: (0, step * j)
    mean = data[..., :3].mean(axis=(0, 1, 2, 3)); std = data[..., :3].std(axis=(0, 1, 2, 3))
    bt = cno.train.DeviceNSFlowBatcher(data, time_indices, start=1, mean=mean, std=std, device='cpu')
    idx = [0, 7, 9, 23, 38]
    time, inp, lab = bt.batch(idx)
    m, s = torch.tensor(mean).view(3, 1, 1), torch.tensor(std).view(3, 1, 1)
    for n, ix in enumerate(idx):
        i, k = ix // len(time_indices), ix % len(time_indices)
        t1, t2 = time_indices[k]
        sample = data[i + 1]
        a = torch.tensor(sample[t1]).permute(2, 0, 1).clone()
        b = torch.tensor(sample[t2, :, :, :3]).permute(2, 0, 1)
        a[:3] = (a[:3] - m) / s
        b = (b - m) / s
        a = torch.cat([a, torch.ones(1, S, S) * ((t2 - t1) / 19.0)], 0)
        assert abs(float(time[n]) - (t2 - t1) / 19.0) < 1e-7
        assert torch.allclose(inp[n], a, atol=1e-6) and torch.allclose(lab[n], b, atol=1e-6)
    assert inp.shape == (5, 7, S, S) and lab.shape == (5, 3, S, S)
    own = cno.train.DeviceNSFlowBatcher(data, time_indices, device='cpu')      # per-file statistics fallback
    assert torch.allclose(own.mean.view(-1), torch.tensor(mean), atol=1e-5) and torch.allclose(own.std.view(-1), torch.tensor(std), atol=1e-5)
