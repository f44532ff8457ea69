"""CPU oracle for the CNO activation hot path  --  TEST INFRASTRUCTURE ONLY.

This file is a restatement (not a copy) of the reference's `impl='ref'` algorithms.  It is only
ever imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs, as the *checker*.  The product path (pregenerating-pde_b200/) never imports it and has no
CPU fallback.

Parity status: PINNED.  The reference ships no golden vectors for this path, so
the oracle is pinned against outputs of the reference itself: oracle/make_golden.py imports the
unmodified reference from /root/reference/CNO_Experiments, runs its `_filtered_lrelu_ref`,
`_upfirdn2d_ref`, `_bias_act_ref` and `FILM.forward` on seeded inputs and stores inputs+outputs
under tests/golden/*.npz; tests/test_oracle_golden.py re-runs this oracle on the stored inputs
and requires agreement to fp32 round-off.

Each function cites the reference lines it follows (paths relative to
/root/reference/CNO_Experiments/).
"""
import math
import numpy as np
import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------------------------
# argument normalisation (torch_utils/ops/upfirdn2d.py:27-58, filtered_lrelu.py:27-44)
# --------------------------------------------------------------------------------------------
def pair(v):
    if isinstance(v, (int, np.integer)):
        return int(v), int(v)
    a, b = v
    return int(a), int(b)


def pad4(padding):
    if isinstance(padding, (int, np.integer)):
        padding = [padding, padding]
    padding = [int(p) for p in padding]
    if len(padding) == 2:
        padding = [padding[0], padding[0], padding[1], padding[1]]
    assert len(padding) == 4
    return tuple(padding)


def filter_wh(f):
    if f is None:
        return 1, 1
    assert f.ndim in (1, 2)
    return int(f.shape[-1]), int(f.shape[0])


# --------------------------------------------------------------------------------------------
# bias_act  (torch_utils/ops/bias_act.py:13-23 table, :83-112 _bias_act_ref)
# --------------------------------------------------------------------------------------------
ACTIVATIONS = {
    # name: (fn(x, alpha), default alpha, default gain, cuda_idx)
    'linear':   (lambda x, a: x,                         0.0, 1.0,            1),
    'relu':     (lambda x, a: F.relu(x),                 0.0, math.sqrt(2.0), 2),
    'lrelu':    (lambda x, a: F.leaky_relu(x, a),        0.2, math.sqrt(2.0), 3),
    'tanh':     (lambda x, a: torch.tanh(x),             0.0, 1.0,            4),
    'sigmoid':  (lambda x, a: torch.sigmoid(x),          0.0, 1.0,            5),
    'elu':      (lambda x, a: F.elu(x),                  0.0, 1.0,            6),
    'selu':     (lambda x, a: F.selu(x),                 0.0, 1.0,            7),
    'softplus': (lambda x, a: F.softplus(x),             0.0, 1.0,            8),
    'swish':    (lambda x, a: torch.sigmoid(x) * x,      0.0, math.sqrt(2.0), 9),
}


def bias_act(x, b=None, dim=1, act='linear', alpha=None, gain=None, clamp=None):
    """y = clamp(act(x + b) * gain).  bias_act.py:83-112."""
    fn, def_alpha, def_gain, _ = ACTIVATIONS[act]
    alpha = float(def_alpha if alpha is None else alpha)
    gain = float(def_gain if gain is None else gain)
    if b is not None:
        shape = [1] * x.ndim
        shape[dim] = -1
        x = x + b.reshape(shape)
    x = fn(x, alpha)
    if gain != 1.0:
        x = x * gain
    if clamp is not None and clamp >= 0:
        x = x.clamp(-clamp, clamp)
    return x


# --------------------------------------------------------------------------------------------
# upfirdn2d  (torch_utils/ops/upfirdn2d.py:159-203 _upfirdn2d_ref)
# --------------------------------------------------------------------------------------------
def upfirdn2d(x, f, up=1, down=1, padding=0, flip_filter=False, gain=1):
    """zero-insert upsample -> pad/crop -> FIR (true convolution unless flip_filter) -> decimate."""
    assert x.ndim == 4
    n, c, h, w = x.shape
    upx, upy = pair(up)
    downx, downy = pair(down)
    px0, px1, py0, py1 = pad4(padding)
    if f is None:
        f = torch.ones([1, 1], dtype=torch.float32)
    # :179-181  zero insertion: sample (i) lands on position i*up
    z = x.new_zeros([n, c, h * upy, w * upx])
    z[:, :, ::upy, ::upx] = x
    # :184-185  positive padding pads, negative padding crops
    z = F.pad(z, [max(px0, 0), max(px1, 0), max(py0, 0), max(py1, 0)])
    z = z[:, :, max(-py0, 0): z.shape[2] - max(-py1, 0), max(-px0, 0): z.shape[3] - max(-px1, 0)]
    # :188-191  gain split evenly over the separable passes, filter cast to x.dtype, flipped for conv
    k = (f * (gain ** (f.ndim / 2))).to(x.dtype)
    if not flip_filter:
        k = k.flip(list(range(k.ndim)))
    # :194-199  depth-wise correlation; separable = horizontal pass then vertical pass
    if k.ndim == 2:
        z = F.conv2d(z, k[None, None].expand(c, 1, -1, -1).contiguous(), groups=c)
    else:
        z = F.conv2d(z, k[None, None, None, :].expand(c, 1, 1, -1).contiguous(), groups=c)
        z = F.conv2d(z, k[None, None, :, None].expand(c, 1, -1, 1).contiguous(), groups=c)
    # :202
    return z[:, :, ::downy, ::downx]


# --------------------------------------------------------------------------------------------
# filtered_lrelu  (torch_utils/ops/filtered_lrelu.py:113-145 _filtered_lrelu_ref)
# --------------------------------------------------------------------------------------------
def filtered_lrelu_out_size(in_hw, fu, fd, up, down, padding):
    """filtered_lrelu.py:133-134 (== filtered_lrelu.cpp:61-69)."""
    fu_w, fu_h = filter_wh(fu)
    fd_w, fd_h = filter_wh(fd)
    px0, px1, py0, py1 = pad4(padding)
    in_h, in_w = in_hw
    out_w = (in_w * up + (px0 + px1) - (fu_w - 1) - (fd_w - 1) + (down - 1)) // down
    out_h = (in_h * up + (py0 + py1) - (fu_h - 1) - (fd_h - 1) + (down - 1)) // down
    return out_h, out_w


def filtered_lrelu_preact(x, fu=None, b=None, up=1, padding=0, gain=math.sqrt(2), flip_filter=False):
    """The pre-activation `v` whose sign the 2-bit mask records (v = upFIR(x+b)*up^2, times gain).
    filtered_lrelu.py:137-138; mask semantics filtered_lrelu.cu:622-635."""
    v = bias_act(x, b)
    v = upfirdn2d(v, fu, up=up, padding=pad4(padding), gain=up ** 2, flip_filter=flip_filter)
    return v * gain


def filtered_lrelu(x, fu=None, fd=None, b=None, up=1, down=1, padding=0, gain=math.sqrt(2),
                   slope=0.2, clamp=None, flip_filter=False):
    """bias -> upFIR(up^2) -> lrelu*gain, clamp -> downFIR -> decimate.  filtered_lrelu.py:137-140."""
    assert x.ndim == 4
    v = bias_act(x, b)
    v = upfirdn2d(v, fu, up=up, padding=pad4(padding), gain=up ** 2, flip_filter=flip_filter)
    v = bias_act(v, act='lrelu', alpha=slope, gain=gain, clamp=clamp)
    y = upfirdn2d(v, fd, down=down, flip_filter=flip_filter)
    oh, ow = filtered_lrelu_out_size(x.shape[2:], fu, fd, up, down, padding)
    assert tuple(y.shape[2:]) == (oh, ow)
    return y


def sign_mask_shape(out_hw, fd, down):
    """[sh, sw_bytes, sw_valid]: filtered_lrelu.cpp:79-86."""
    fd_w, fd_h = filter_wh(fd)
    oh, ow = out_hw
    sh = oh * down - (down - 1) + (fd_h - 1)
    sw_valid = ow * down - (down - 1) + (fd_w - 1)
    sw = (sw_valid + 15) & ~15
    return sh, sw // 4, sw_valid


def sign_mask_from_preact(v, slope, gain, clamp, sh, sw_valid):
    """2-bit codes per element of the up-res pre-activation: 1 = negative, 2 = clamped (clamp
    overrides sign), 0 = pass.  filtered_lrelu.cu:622-635.  `v` already includes `gain`."""
    code = (v < 0).to(torch.uint8)
    if clamp is not None:
        act = torch.where(v < 0, v * slope, v)
        code = torch.where(act.abs() > clamp, torch.full_like(code, 2), code)
    return code[:, :, :sh, :sw_valid]


def pack_sign_mask(code, sw_bytes):
    """element x -> byte x>>2, bits (x&3)*2.  filtered_lrelu.cu sign packing, SURVEY §7.3-4."""
    n, c, sh, swv = code.shape
    full = torch.zeros([n, c, sh, sw_bytes * 4], dtype=torch.uint8)
    full[..., :swv] = code
    q = full.reshape(n, c, sh, sw_bytes, 4).to(torch.int32)
    packed = q[..., 0] | (q[..., 1] << 2) | (q[..., 2] << 4) | (q[..., 3] << 6)
    return packed.to(torch.uint8)


def unpack_sign_mask(packed, sw_valid):
    p = packed.to(torch.int32)
    parts = [(p >> (2 * i)) & 3 for i in range(4)]
    full = torch.stack(parts, dim=-1).reshape(*packed.shape[:-1], packed.shape[-1] * 4)
    return full[..., :sw_valid].to(torch.uint8)


# --------------------------------------------------------------------------------------------
# CIN / FILM  (CNO_timeModule_CIN.py:27-70, instance-norm branch nl_dim == [2,3])
# --------------------------------------------------------------------------------------------
def cin(x, gamma, beta, scale, shift, eps=1e-5):
    """InstanceNorm2d(affine) followed by per-(n,c) scale/shift.
    x [N,C,H,W]; gamma, beta [C] (norm.weight / norm.bias); scale, shift [N,C]
    (= lat2scale(inp2lat_sacale(t)), lat2bias(inp2lat_bias(t)))."""
    xn = F.instance_norm(x, weight=gamma, bias=beta, eps=eps)
    return xn * scale[:, :, None, None] + shift[:, :, None, None]


def film_forward(x, time, w):
    """Full FILM.forward (:60-70) from a dict of its parameters (reference state_dict names)."""
    t = time.reshape(-1, 1).to(x.dtype)
    lat_s = F.linear(t, w['inp2lat_sacale.weight'], w['inp2lat_sacale.bias'])
    scale = F.linear(lat_s, w['lat2scale.weight'], w['lat2scale.bias'])
    lat_b = F.linear(t, w['inp2lat_bias.weight'], w['inp2lat_bias.bias'])
    shift = F.linear(lat_b, w['lat2bias.weight'], w['lat2bias.bias'])
    return cin(x, w['norm.weight'], w['norm.bias'], scale, shift)


# --------------------------------------------------------------------------------------------
# LReLu filter design + padding  (training/filtered_networks.py:231-262, :280-291)
# --------------------------------------------------------------------------------------------
def design_lowpass_filter(numtaps, cutoff, width, fs):
    import scipy.signal
    if numtaps == 1:
        return None
    return torch.as_tensor(scipy.signal.firwin(numtaps=numtaps, cutoff=cutoff, width=width, fs=fs),
                           dtype=torch.float32)


def lrelu_config(in_size, out_size, filter_size=6, lrelu_upsampling=2):
    """The (fu, fd, up, down, padding) that CNOBlock's LReLu builds for in_size -> out_size
    (CNO_timeModule_CIN.py:285-298 passes cutoff = size/2.0001, half_width = 0.8*size - cutoff,
    sampling rate = size)."""
    in_cut, out_cut = in_size / 2.0001, out_size / 2.0001
    in_hw, out_hw = 0.8 * in_size - in_cut, 0.8 * out_size - out_cut
    tmp_rate = max(in_size, out_size) * lrelu_upsampling
    up = int(np.rint(tmp_rate / in_size))
    down = int(np.rint(tmp_rate / out_size))
    up_taps = filter_size * up if up > 1 else 1
    down_taps = filter_size * down if down > 1 else 1
    fu = design_lowpass_filter(up_taps, in_cut, in_hw * 2, tmp_rate)
    fd = design_lowpass_filter(down_taps, out_cut, out_hw * 2, tmp_rate)
    pad_total = (out_size - 1) * down + 1 - in_size * up + up_taps + down_taps - 2
    lo = (pad_total + up) // 2
    hi = pad_total - lo
    return dict(fu=fu, fd=fd, up=up, down=down, padding=[lo, hi, lo, hi])
