"""Fine-tuning wrappers of the CNO foundation model: the callers either side of the hot path in CNO_FineTune.py.

`FT_Lift`, `FT_Project`, `initialize_FT` mirror test_and_fine_tune_utils/fine_tune_lift.py:12-108 (attribute names kept, so
state_dict keys match), `finetune_param_groups` the three learning-rate groups of CNO_timeModule_CIN.py:983-994.
`wide_finetune_config` is BASELINE.json configs[4]: the foundation-model shape at 256 x 256 with channel_multiplier 64
(features 32 ... 512; SURVEY 8d "Wide FT").  Nothing here computes on the GPU by itself: the modules are 1x1 convolutions
around the model's own lift / project blocks, whose activations are the package's fused kernels.
"""
import torch
from torch import nn

from . import cno_model


class FT_Lift(nn.Module):
    """1x1 convolutions in front of the pretrained lift (fine_tune_lift.py:12-46): new_in -> 128 -> old_in when the channel
    count changed, else one square 1x1 convolution."""

    def __init__(self, in_channels, latent_channels, model, pre_lift=False):
        super().__init__()
        self.latent_channels, self.pre_lift = latent_channels, pre_lift
        if pre_lift:
            self.linear1 = nn.Conv2d(in_channels, 128, kernel_size=1, padding=0)
            self.linear2 = nn.Conv2d(128, latent_channels, kernel_size=1, padding=0)
        else:
            self.linear1 = nn.Conv2d(in_channels, in_channels, kernel_size=1, padding=0)
            self.linear2 = nn.Identity()
        self.lift = model.lift

    def forward(self, x, timestep):
        return self.lift(self.linear2(self.linear1(x)), timestep)


class FT_Project(nn.Module):
    """Either a fresh projection block for a new output channel count -- whose out_size is the constant 128 of
    fine_tune_lift.py:58-64, so a 256^2 input leaves at 128^2 through one extra 256 -> 128 down-sampling activation -- or the
    pretrained projection followed by two 1x1 convolutions (fine_tune_lift.py:49-84)."""

    def __init__(self, new_out_dim, model, init_project=False):
        super().__init__()
        if init_project:
            self.in_channels = model.encoder_features[0] + model.decoder_features_out[-1]
            self.project = cno_model.LiftProjectBlock(self.in_channels, new_out_dim, model.decoder_sizes[-1], 128, latent_dim=64)
            self.post_linear1 = nn.Identity()
            self.post_linear2 = nn.Identity()
        else:
            self.project = model.project
            self.post_linear1 = nn.Conv2d(new_out_dim, 128, kernel_size=1, padding=0)
            self.post_linear2 = nn.Conv2d(128, new_out_dim, kernel_size=1, padding=0)

    def forward(self, x, timestep):
        return self.post_linear2(self.post_linear1(self.project(x, timestep)))


def initialize_FT(model, old_in_dim, new_in_dim, new_out_dim, old_out_dim):
    """swap lift / project of a pretrained CNO_time for the fine-tuning wrappers (fine_tune_lift.py:86-108)"""
    model.lift = FT_Lift(new_in_dim, old_in_dim, model, pre_lift=old_in_dim != new_in_dim)
    model.project = FT_Project(new_out_dim, model, init_project=new_out_dim != old_out_dim)
    return model


def finetune_param_groups(model, lr, lr_emb, lr_norm):
    """AdamW groups of a fine-tuning run (CNO_timeModule_CIN.py:983-994): trunk at lr, lift / project at lr_emb, the
    time-conditioned norms outside lift / project at lr_norm."""
    named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
    lp = lambda n: ('project' in n) or ('lift' in n)
    cond = lambda n: 'in_norm_conditiner' in n
    g1 = [p for n, p in named if not lp(n) and not cond(n)]
    g2 = [p for n, p in named if lp(n) and not cond(n)]
    g3 = [p for n, p in named if cond(n) and not lp(n)]
    return [{'params': g1, 'lr': lr}, {'params': g2, 'lr': lr_emb}, {'params': g3, 'lr': lr_norm}]


def wide_finetune_config():
    """BASELINE.json configs[4]: foundation model of CNO_FineTune.py:62-63 (in_dim 5, out_dim 4, N_layers 4, N_res 8,
    N_res_neck 8) built at in_size 256 with channel_multiplier 64 -> features [32, 64, 128, 256, 512], 67.7 M parameters;
    fine-tuned to 4 input / 3 output channels (CNO_FineTune.py:41-42,90-94) with the default learning rates (:27-31)."""
    return dict(model=dict(in_dim=5, in_size=256, N_layers=4, N_res=8, N_res_neck=8, channel_multiplier=64, out_dim=4, is_time=1,
                           nl_dim=[2, 3], lr=0.00005, weight_decay=1e-6),
                ft=dict(old_in_dim=5, new_in_dim=4, new_out_dim=3, old_out_dim=4),
                lrs=dict(lr=0.00005, lr_emb=0.0005, lr_norm=0.00125), in_size=256, out_size=128, in_dim=4, out_dim=3)
