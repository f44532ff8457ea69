"""Attach the B200 kernels to an (unmodified) checkout of the reference's CNO_Experiments.

    import pregenerating_pde_b200 as cno
    cno.patch_reference('/path/to/CNO_Experiments')     # before or after importing the reference modules

After this, `torch_utils.ops.{filtered_lrelu,upfirdn2d,bias_act}` resolve their `_plugin` through our
loader (the reference's Python wrappers run unchanged on the C-ABI library) and `FILM.forward` is the
fused CIN.  Parameter names / state_dict keys are untouched (SURVEY §5 checkpoint constraint)."""
import importlib
import sys

import torch

from .torch_utils import custom_ops as _our_custom_ops
from . import cin as _cin

_saved = {}


def patch_reference(cno_root=None, patch_film=True):
    if cno_root is not None and cno_root not in sys.path:
        sys.path.insert(0, cno_root)
    ref_custom_ops = importlib.import_module('torch_utils.custom_ops')
    if 'get_plugin' not in _saved:
        _saved['get_plugin'] = ref_custom_ops.get_plugin
    ref_custom_ops.get_plugin = _our_custom_ops.get_plugin
    for name in ('bias_act', 'upfirdn2d', 'filtered_lrelu'):
        mod = importlib.import_module(f'torch_utils.ops.{name}')
        mod._plugin = None            # force re-resolution through the patched loader
    if patch_film:
        try:
            film_mod = importlib.import_module('CNO_timeModule_CIN')
        except Exception:             # model module not importable (missing lightning etc.): ops only
            film_mod = None
        if film_mod is not None:
            if 'film_forward' not in _saved:
                _saved['film_forward'] = film_mod.FILM.forward
            film_mod.FILM.forward = _cin.film_forward
            if hasattr(film_mod, 'LReLu'):
                _cin.register_lrelu(film_mod.LReLu)      # training/filtered_networks.py:267-276 is the call film_lrelu replaces
            if hasattr(film_mod, 'CNOBlock'):
                if 'cno_block_forward' not in _saved:
                    _saved['cno_block_forward'] = film_mod.CNOBlock.forward
                film_mod.CNOBlock.forward = _cno_block_forward
            if hasattr(film_mod, 'ResidualBlock'):
                if 'residual_forward' not in _saved:
                    _saved['residual_forward'] = film_mod.ResidualBlock.forward
                film_mod.ResidualBlock.forward = _residual_forward
    return True


def _cno_block_forward(self, x, time):
    """CNOBlock.forward (CNO_timeModule_CIN.py:329-335) with the CIN folded into the activation's prologue (cin.film_lrelu)."""
    timed = self.is_time == 1 or self.is_time is True
    if not (timed and isinstance(self.batch_norm, torch.nn.Identity)):
        return _saved['cno_block_forward'](self, x, time)
    return _cin.film_lrelu(self.in_norm_conditiner, self.activation, _cin.conv_before_film(self.convolution, self.in_norm_conditiner, x), time)


def _residual_forward(self, x, time):
    """ResidualBlock.forward (CNO_timeModule_CIN.py:496-510) with the first CIN folded into the activation's prologue and the closing
    `x + out` folded into the second CIN's kernel.
    Any flavour other than time conditioning with Identity batch norms runs the reference's own forward."""
    timed = self.is_time == 1 or self.is_time is True
    if not (timed and isinstance(self.batch_norm1, torch.nn.Identity) and isinstance(self.batch_norm2, torch.nn.Identity)):
        return _saved['residual_forward'](self, x, time)
    out = _cin.film_lrelu(self.in_norm_conditiner1, self.activation, _cin.conv_before_film(self.convolution1, self.in_norm_conditiner1, x), time)
    out = _cin.conv_before_film(self.convolution2, self.in_norm_conditiner2, out)
    return self.in_norm_conditiner2(out, time, residual=x)


def unpatch_reference():
    if 'get_plugin' in _saved:
        importlib.import_module('torch_utils.custom_ops').get_plugin = _saved.pop('get_plugin')
        for name in ('bias_act', 'upfirdn2d', 'filtered_lrelu'):
            importlib.import_module(f'torch_utils.ops.{name}')._plugin = None
    if 'film_forward' in _saved:
        importlib.import_module('CNO_timeModule_CIN').FILM.forward = _saved.pop('film_forward')
    if 'cno_block_forward' in _saved:
        importlib.import_module('CNO_timeModule_CIN').CNOBlock.forward = _saved.pop('cno_block_forward')
    if 'residual_forward' in _saved:
        importlib.import_module('CNO_timeModule_CIN').ResidualBlock.forward = _saved.pop('residual_forward')
