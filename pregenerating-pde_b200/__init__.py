"""B200-native CNO activation stack: filtered_lrelu / upfirdn2d / bias_act / CIN.

Mirrors the reference's `torch_utils.ops` API (CNO_Experiments/torch_utils/ops/*.py) on top of
hand-written sm_100a kernels reached through the C ABI of libcno_b200.so (include/cno_b200.h).
There is no CPU fallback in this package: `impl='ref'` is the pure-torch definition kept for API
parity (as in the reference), `impl='cuda'` on a CUDA tensor requires the compiled library.
"""
from . import _lib                                   # noqa: F401
from .torch_utils import custom_ops                  # noqa: F401
from .torch_utils.ops import bias_act, upfirdn2d, filtered_lrelu   # noqa: F401
from . import cin                                    # noqa: F401
from . import layout, optim                          # noqa: F401
from . import cno_model, train, finetune             # noqa: F401
from .patch import patch_reference, unpatch_reference              # noqa: F401

__all__ = ['bias_act', 'upfirdn2d', 'filtered_lrelu', 'cin', 'custom_ops', 'patch_reference', 'unpatch_reference']
