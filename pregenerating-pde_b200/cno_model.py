"""Host-side mirror of the CALLERS of the hot path: `LReLu` (training/filtered_networks.py:199-311), `FILM`,
`CNOBlock`, `LiftProjectBlock`, `ResidualBlock` and the `CNO_time` network
(CNO_timeModule_CIN.py:27-70, :227-335, :338-387, :393-510, :514-884).

Why it exists: the reference model imports pytorch_lightning / matplotlib / h5py / netCDF4 and lives in
/root/reference, neither of which is present on the GPU box, so the CNO training benchmark (BASELINE.json
configs[2..4]) needs a definition that travels with this repo.  It is NOT a re-implementation of the
model's maths beyond what the reference modules do: every module keeps the reference's attribute names,
so `state_dict()` keys are identical (tests/test_cno_model.py loads a reference-initialised state dict and
compares forward + gradients against the unmodified reference model on CPU), and the hot ops are the
package's `filtered_lrelu` and fused CIN.  Only the instance-norm conditioning flavour (`nl_dim=[2,3]`,
the TrainCNO_time_L.py configuration) plus the un-normalised lift/project blocks are covered; the ViT
bottleneck, BatchNorm / LayerNorm flavours and the alternative activations are out of scope.
"""
import numpy as np
import scipy.signal
import scipy.special
import torch
import torch.nn.functional as F
from torch import nn

from .torch_utils.ops import filtered_lrelu
from . import cin as _cin


@_cin.register_lrelu
class LReLu(nn.Module):
    """Alias-free activation: up-FIR -> leaky ReLU -> down-FIR (filtered_networks.py:199-278)."""

    def __init__(self, in_channels, out_channels, in_size, out_size, in_sampling_rate, out_sampling_rate,
                 in_cutoff, out_cutoff, in_half_width, out_half_width, filter_size=6, lrelu_upsampling=2,
                 is_critically_sampled=False, use_radial_filters=False, impl='cuda'):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.in_size = np.broadcast_to(np.asarray(in_size), [2])
        self.out_size = np.broadcast_to(np.asarray(out_size), [2])
        self.tmp_sampling_rate = max(in_sampling_rate, out_sampling_rate) * lrelu_upsampling
        self.impl = impl
        self.bias = nn.Parameter(torch.zeros([out_channels]))
        self.up_factor = int(np.rint(self.tmp_sampling_rate / in_sampling_rate))
        self.up_taps = filter_size * self.up_factor if self.up_factor > 1 else 1
        self.register_buffer('up_filter', self.design_lowpass_filter(self.up_taps, in_cutoff, in_half_width * 2, self.tmp_sampling_rate))
        self.down_factor = int(np.rint(self.tmp_sampling_rate / out_sampling_rate))
        self.down_taps = filter_size * self.down_factor if self.down_factor > 1 else 1
        self.down_radial = bool(use_radial_filters) and not is_critically_sampled            # filtered_networks.py:245
        self.register_buffer('down_filter', self.design_lowpass_filter(self.down_taps, out_cutoff, out_half_width * 2, self.tmp_sampling_rate,
                                                                       radial=self.down_radial))
        # padding of the up-sampled signal (filtered_networks.py:254-262)
        total = (self.out_size - 1) * self.down_factor + 1 - self.in_size * self.up_factor + self.up_taps + self.down_taps - 2
        lo = (total + self.up_factor) // 2
        hi = total - lo
        self.padding = [int(lo[0]), int(hi[0]), int(lo[1]), int(hi[1])]

    @staticmethod
    def design_lowpass_filter(numtaps, cutoff, width, fs, radial=False):
        """Kaiser low-pass taps (separable) or the radially symmetric jinc filter [numtaps, numtaps]
        (filtered_networks.py:280-301); a 2-D down filter takes the generic composition instead of the fused kernel."""
        if numtaps == 1:
            return None
        if not radial:
            return torch.as_tensor(scipy.signal.firwin(numtaps=numtaps, cutoff=cutoff, width=width, fs=fs), dtype=torch.float32)
        pos = (np.arange(numtaps) - (numtaps - 1) / 2) / fs
        rad = np.hypot(*np.meshgrid(pos, pos))
        f = scipy.special.j1(2 * cutoff * (np.pi * rad)) / (np.pi * rad)
        win = np.kaiser(numtaps, scipy.signal.kaiser_beta(scipy.signal.kaiser_atten(numtaps, width / (fs / 2))))
        f *= np.outer(win, win)
        return torch.as_tensor(f / np.sum(f), dtype=torch.float32)

    def forward(self, x):
        y = filtered_lrelu.filtered_lrelu(x=x, fu=self.up_filter, fd=self.down_filter, b=self.bias.to(x.dtype), up=self.up_factor,
                                          down=self.down_factor, padding=self.padding, gain=np.sqrt(2), slope=0.2, clamp=None,
                                          impl=self.impl)
        assert y.shape[1:] == (self.out_channels, int(self.out_size[1]), int(self.out_size[0]))
        return y


class LReLu_torch(nn.Module):
    """The reference's torch-only activation (training/filtered_networks.py:349-388, `activation='cno_lrelu_torch'` of
    CNOBlock / ResidualBlock, CNO_timeModule_CIN.py:299-305, :461-467): bicubic antialiased interpolation to 2x, LeakyReLU
    (default slope 0.01), back to the input size, on to the output size, bias added last.  Plain torch ops (no kernel of this
    package is involved); kept so that checkpoints and configurations that select it load and run."""

    def __init__(self, in_channels, out_channels, in_size, out_size, in_sampling_rate=None, out_sampling_rate=None):
        super().__init__()
        assert in_channels == out_channels
        self.activation = nn.LeakyReLU()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.in_size, self.out_size = in_size, out_size
        self.in_sampling_rate, self.out_sampling_rate = in_sampling_rate, out_sampling_rate
        self.bias = nn.Parameter(torch.zeros([out_channels]))

    def forward(self, x):
        x = F.interpolate(x, size=2 * self.in_size, mode='bicubic', antialias=True)
        x = self.activation(x)
        x = F.interpolate(x, size=self.in_size, mode='bicubic', antialias=True)
        x = F.interpolate(x, size=self.out_size, mode='bicubic', antialias=True)
        return x + self.bias.view(1, -1, 1, 1)


def _make_activation(kind, channels, in_size, out_size, cutoff_den, half_width_mult, filter_size, lrelu_upsampling, impl):
    if kind == 'cno_lrelu':
        return LReLu(channels, channels, in_size, out_size, filter_size=filter_size, lrelu_upsampling=lrelu_upsampling,
                     is_critically_sampled=(cutoff_den == 2.0), impl=impl, **_filter_args(in_size, out_size, cutoff_den, half_width_mult))
    if kind == 'cno_lrelu_torch':
        return LReLu_torch(channels, channels, in_size, out_size, in_size, out_size)
    raise ValueError('Please specify different activation function')      # CNO_timeModule_CIN.py:315


class FILM(nn.Module):
    """Time-conditioned instance norm (CNO_timeModule_CIN.py:27-70); forward = the fused CIN kernel."""

    def __init__(self, channels, dim=(2, 3), s=128, intermediate=128):
        super().__init__()
        assert list(dim) == [2, 3], 'only the instance-norm flavour (nl_dim "23") is mirrored'
        self.channels, self.s = channels, s
        self.inp2lat_sacale = nn.Linear(1, intermediate)
        self.lat2scale = nn.Linear(intermediate, channels)
        self.inp2lat_bias = nn.Linear(1, intermediate)
        self.lat2bias = nn.Linear(intermediate, channels)
        with torch.no_grad():                       # identity conditioning at init (:43-49)
            self.inp2lat_sacale.weight.zero_(); self.lat2scale.weight.zero_(); self.lat2scale.bias.fill_(1)
            self.inp2lat_bias.weight.zero_(); self.lat2bias.weight.zero_(); self.lat2bias.bias.zero_()
        self.norm = nn.InstanceNorm2d(channels, affine=True)

    forward = _cin.film_forward


def _filter_args(in_size, out_size, cutoff_den, half_width_mult):
    return dict(in_sampling_rate=in_size, out_sampling_rate=out_size, in_cutoff=in_size / cutoff_den,
                out_cutoff=out_size / cutoff_den, in_half_width=half_width_mult * in_size - in_size / cutoff_den,
                out_half_width=half_width_mult * out_size - out_size / cutoff_den)


class CNOBlock(nn.Module):
    """conv3x3 -> (CIN) -> LReLu; D / U / I by the in/out sizes (CNO_timeModule_CIN.py:227-335)."""

    def __init__(self, in_channels, out_channels, in_size, out_size, cutoff_den=2.0001, conv_kernel=3, filter_size=6,
                 lrelu_upsampling=2, half_width_mult=0.8, is_time=True, nl_dim=(2, 3), impl='cuda', activation='cno_lrelu'):
        super().__init__()
        self.in_channels, self.out_channels, self.in_size, self.out_size = in_channels, out_channels, in_size, out_size
        self.convolution = nn.Conv2d(in_channels, out_channels, conv_kernel, padding=(conv_kernel - 1) // 2)
        self.batch_norm = nn.Identity()
        self.activation = _make_activation(activation, out_channels, in_size, out_size, cutoff_den, half_width_mult, filter_size,
                                           lrelu_upsampling, impl)
        self.is_time = bool(is_time)
        if self.is_time:
            self.in_norm_conditiner = FILM(out_channels, dim=nl_dim, s=in_size)

    def forward(self, x, time):
        if self.is_time:
            x = _cin.conv_before_film(self.convolution, self.in_norm_conditiner, x)       # the bias would vanish in the instance norm
            return _cin.film_lrelu(self.in_norm_conditiner, self.activation, x, time)     # CIN folded into the activation's prologue
        return self.activation(self.convolution(x))


class LiftProjectBlock(nn.Module):
    """CNOBlock (no norm) -> conv3x3 (CNO_timeModule_CIN.py:338-387)."""

    def __init__(self, in_channels, out_channels, in_size, out_size, latent_dim=64, conv_kernel=3, **kw):
        super().__init__()
        self.out_channels = out_channels
        self.inter_CNOBlock = CNOBlock(in_channels, latent_dim, in_size, out_size, conv_kernel=conv_kernel, is_time=False, **kw)
        self.convolution = nn.Conv2d(latent_dim, out_channels, conv_kernel, stride=1, padding=(conv_kernel - 1) // 2)

    def forward(self, x, time):
        return _cin.conv_nhwc(self.convolution, self.inter_CNOBlock(x, time))


class ResidualBlock(nn.Module):
    """x + CIN(conv(LReLu(CIN(conv(x))))) (CNO_timeModule_CIN.py:393-510)."""

    def __init__(self, channels, size, cutoff_den=2.0001, conv_kernel=3, filter_size=6, lrelu_upsampling=2,
                 half_width_mult=0.8, is_time=True, nl_dim=(2, 3), impl='cuda', activation='cno_lrelu'):
        super().__init__()
        self.channels, self.size = channels, size
        pad = (conv_kernel - 1) // 2
        self.convolution1 = nn.Conv2d(channels, channels, conv_kernel, stride=1, padding=pad)
        self.convolution2 = nn.Conv2d(channels, channels, conv_kernel, stride=1, padding=pad)
        self.batch_norm1, self.batch_norm2 = nn.Identity(), nn.Identity()
        self.activation = _make_activation(activation, channels, size, size, cutoff_den, half_width_mult, filter_size, lrelu_upsampling, impl)
        self.is_time = bool(is_time)
        if self.is_time:
            self.in_norm_conditiner1 = FILM(channels, dim=nl_dim, s=size)
            self.in_norm_conditiner2 = FILM(channels, dim=nl_dim, s=size)

    def forward(self, x, time):
        if self.is_time:
            out = _cin.conv_before_film(self.convolution1, self.in_norm_conditiner1, x)
            out = _cin.film_lrelu(self.in_norm_conditiner1, self.activation, out, time)
            out = _cin.conv_before_film(self.convolution2, self.in_norm_conditiner2, out)
            return self.in_norm_conditiner2(out, time, residual=x)      # x + CIN(out) in the CIN kernel's single pass
        return x + self.convolution2(self.activation(self.convolution1(x)))


class CNO_time(nn.Module):
    """The time-conditioned CNO U-Net (CNO_timeModule_CIN.py:514-884) as a plain nn.Module (the reference derives
    from LightningModule; loss / optimiser live in `training_loss` / `configure_optimizer` below)."""

    def __init__(self, in_dim, in_size, N_layers, N_res=1, N_res_neck=6, channel_multiplier=32, conv_kernel=3,
                 cutoff_den=2.0001, filter_size=6, lrelu_upsampling=2, half_width_mult=0.8, out_dim=1,
                 latent_lift_proj_dim=64, add_inv=True, is_time=1, nl_dim=(2, 3), lr=0.0005, weight_decay=1e-6, impl='cuda',
                 project_out_size=None, activation='cno_lrelu'):
        super().__init__()
        self.N_layers, self.N_res, self.N_res_neck = int(N_layers), int(N_res), int(N_res_neck)
        self.in_dim, self.out_dim, self.add_inv = in_dim, out_dim, add_inv
        self.lr, self.weight_decay = lr, weight_decay
        L = self.N_layers
        enc = [channel_multiplier // 2] + [2 ** i * channel_multiplier for i in range(L)]            # :564-566
        dec_in = enc[1:][::-1]
        dec_out = enc[:-1][::-1]
        for i in range(1, L):
            dec_in[i] = 2 * dec_in[i]                                                                 # skip concatenation
        inv = dec_in + [enc[0] + dec_out[-1]]                                                         # :598-599
        self.encoder_features, self.decoder_features_in, self.decoder_features_out, self.inv_features = enc, dec_in, dec_out, inv
        self.encoder_sizes = [in_size // 2 ** i for i in range(L + 1)]
        self.decoder_sizes = [in_size // 2 ** (L - i) for i in range(L + 1)]
        kw = dict(cutoff_den=cutoff_den, conv_kernel=conv_kernel, filter_size=filter_size, lrelu_upsampling=lrelu_upsampling,
                  half_width_mult=half_width_mult, impl=impl, activation=activation)     # 'cno_lrelu' | 'cno_lrelu_torch' (:534)
        tkw = dict(kw, is_time=is_time, nl_dim=nl_dim)
        es, ds = self.encoder_sizes, self.decoder_sizes
        self.lift = LiftProjectBlock(in_dim, enc[0], in_size, es[0], latent_dim=latent_lift_proj_dim, **kw)
        self.project = LiftProjectBlock(enc[0] + dec_out[-1], out_dim, ds[-1], project_out_size or in_size,
                                        latent_dim=latent_lift_proj_dim, **kw)
        self.encoder = nn.ModuleList([CNOBlock(enc[i], enc[i + 1], es[i], es[i + 1], **tkw) for i in range(L)])
        self.ED_expansion = nn.ModuleList([CNOBlock(enc[i], enc[i], es[i], ds[L - i], **tkw) for i in range(L + 1)])
        self.decoder = nn.ModuleList([CNOBlock(dec_in[i], dec_out[i], ds[i], ds[i + 1], **tkw) for i in range(L)])
        self.decoder_inv = nn.ModuleList([CNOBlock(inv[i], inv[i], ds[i], ds[i], **tkw) for i in range(L + 1)])
        res = [ResidualBlock(enc[l], es[l], **tkw) for l in range(L) for _ in range(self.N_res)]
        res += [ResidualBlock(enc[L], es[L], **tkw) for _ in range(self.N_res_neck)]
        self.res_nets = nn.Sequential(*res)
        self.transformer = nn.Identity()

    def forward(self, x, time):
        L = self.N_layers
        x = self.lift(x, time)
        skip = []
        for i in range(L):
            for j in range(self.N_res):
                x = self.res_nets[i * self.N_res + j](x, time)
            skip.append(x)
            x = self.encoder[i](x, time)
        for j in range(self.N_res_neck):
            x = self.res_nets[-j - 1](x, time)
        for i in range(L):
            if i == 0:
                x = self.ED_expansion[L](x, time)
            else:
                x = torch.cat((x, self.ED_expansion[L - i](skip[-i], time)), 1)
            if self.add_inv:
                x = self.decoder_inv[i](x, time)
            x = self.decoder[i](x, time)
        x = torch.cat((x, self.ED_expansion[0](skip[0], time)), 1)
        return self.project(x, time)

    @staticmethod
    def training_loss(pred, target):
        """relative L1 (CNO_timeModule_CIN.py:939)."""
        return nn.functional.l1_loss(target, pred) / nn.functional.l1_loss(torch.zeros_like(target), target)

    def configure_optimizer(self):
        """AdamW over all parameters (CNO_timeModule_CIN.py:996)."""
        return torch.optim.AdamW(self.parameters(), lr=self.lr, weight_decay=self.weight_decay)


def train_config_L():
    """Constructor arguments of TrainCNO_time_L.py:51-72,167-195 (FPO: 7 input channels incl. lead time, 3 outputs)."""
    return dict(in_dim=7, in_size=128, N_layers=4, N_res=8, N_res_neck=8, channel_multiplier=32, out_dim=3, is_time=1,
                nl_dim=[2, 3], lr=0.00005, weight_decay=1e-6)


def lrelu_for(channels, in_size, out_size, cutoff_den=2.0001, half_width_mult=0.8, **kw):
    """The LReLu a CNOBlock(in_size -> out_size) builds (CNO_timeModule_CIN.py:285-298): I / D / U configuration."""
    return LReLu(channels, channels, in_size, out_size, **_filter_args(in_size, out_size, cutoff_den, half_width_mult), **kw)
