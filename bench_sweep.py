#!/usr/bin/env python
"""Per-op roofline sweep (BASELINE.json configs[1]): filtered_lrelu / upfirdn2d / bias_act / CIN at the CNO U-Net levels,
fp32 + bf16, forward and backward, on ONE B200.

    python bench_sweep.py [--quick] [--out profiles/sweep_rNN.json]

Every row is timed as a CUDA graph of `reps` back-to-back launches over a ROTATING pool of buffers (pool >= 384 MB where the
tensors allow it, so successive launches miss the 126 MB L2), replayed 5 times, CUDA events around each replay, best replay
reported.  Graph replay removes the Python / ctypes launch cost that would otherwise dominate the small levels.
GB/s = algorithmic bytes / time; frac = GB/s / measured HBM peak (MEASURED_PEAKS.json).
"""
import argparse, json, math, os, sys
ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import torch
import pregenerating_pde_b200 as cno
from pregenerating_pde_b200 import cno_model

DEV = 'cuda'
POOL_BYTES = 384e6


def ref_plugin(name):
    """the reference's own CUDA plugin recompiled for sm_100a (oracle/build_ref_plugins.py), or None.
    Baseline leg only: it is timed beside our kernel and never feeds a product path."""
    if name not in _REF:
        try:
            sys.path.insert(0, os.path.join(ROOT, 'oracle'))
            import build_ref_plugins
            _REF[name] = build_ref_plugins.load(name)
        except Exception as ex:                      # missing build, ABI mismatch, ...: the column stays empty
            print(f'reference plugin {name} unavailable: {ex}', file=sys.stderr)
            _REF[name] = None
    return _REF[name]


_REF = {}


def peak():
    try:
        return json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs']
    except Exception:
        return 6650.0


def time_graph(fn, nrot, reps):
    """fn(i) launches the op on buffer set i % nrot.  Returns best microseconds per launch."""
    for i in range(min(nrot, 3)):
        fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(reps):
            fn(i)
    best = float('inf')
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3 / reps)
    del g
    return best


def rot_count(bytes_per_call):
    return int(max(2, min(48, math.ceil(POOL_BYTES / max(bytes_per_call, 1)))))


def row(op, shape, dtype, direction, us, nbytes, note=''):
    gbs = nbytes / us / 1e3
    return dict(op=op, shape=shape, dtype=str(dtype).replace('torch.', ''), dir=direction, us=round(us, 2),
                algorithmic_bytes=int(nbytes), gbs=round(gbs, 1), frac_hbm=round(gbs / peak(), 4), note=note)


def sweep_filtered_lrelu(rows, N, C, s_in, s_out, dtype, reps):
    act = cno_model.lrelu_for(C, s_in, s_out).to(DEV)
    plug = cno.custom_ops.get_plugin('filtered_lrelu_plugin')
    up, down = act.up_factor, act.down_factor
    px0, px1, py0, py1 = act.padding
    fu, fd = act.up_filter, act.down_filter
    tu, td = fu.numel(), fd.numel()
    es = torch.empty(0, dtype=dtype).element_size()
    sh = s_out * down - (down - 1) + (td - 1)
    swb = ((sh + 15) // 16 * 16) // 4
    nbytes = N * C * (s_in * s_in * es + s_out * s_out * es + sh * swb) + C * es
    nrot = rot_count(nbytes)
    xs = [torch.randn(N, C, s_in, s_in, device=DEV, dtype=dtype) for _ in range(nrot)]
    dys = [torch.randn(N, C, s_out, s_out, device=DEV, dtype=dtype) for _ in range(nrot)]
    b = torch.randn(C, device=DEV, dtype=dtype); zb = torch.zeros_like(b)
    e = torch.empty(0)
    gain, slope = math.sqrt(2), 0.2
    y, so, rc = plug.filtered_lrelu(xs[0], fu, fd, b, e, up, down, px0, px1, py0, py1, 0, 0, gain, slope, float('inf'), False, True)
    assert rc == 0, 'fused kernel refused a CNO configuration'
    masks = [so] + [torch.empty_like(so) for _ in range(nrot - 1)]
    for i in range(1, nrot):          # real masks for every rotation slot
        masks[i] = plug.filtered_lrelu(xs[i], fu, fd, b, e, up, down, px0, px1, py0, py1, 0, 0, gain, slope, float('inf'), False, True)[1]
    pp = ((tu - 1) + (td - 1) - px0, s_in * up - s_out * down + px0 - (up - 1), (tu - 1) + (td - 1) - py0, s_in * up - s_out * down + py0 - (up - 1))
    gg = gain * up * up / (down * down)
    fwd = lambda i: plug.filtered_lrelu(xs[i % nrot], fu, fd, b, e, up, down, px0, px1, py0, py1, 0, 0, gain, slope, float('inf'), False, True)
    bwd = lambda i: plug.filtered_lrelu(dys[i % nrot], fd, fu, zb, masks[i % nrot], down, up, pp[0], pp[1], pp[2], pp[3],
                                        px0 - (tu - 1), py0 - (tu - 1), gg, slope, float('inf'), True, False, signs_only=True)
    cfgname = 'I' if s_in == s_out else ('D' if s_out < s_in else 'U')
    shape = f'[{N},{C},{s_in}->{s_out}] {cfgname}'
    rf, rb = row('filtered_lrelu', shape, dtype, 'fwd', time_graph(fwd, nrot, reps), nbytes), None
    rows.append(rf)
    rb = row('filtered_lrelu', shape, dtype, 'bwd', time_graph(bwd, nrot, reps), nbytes)
    rows.append(rb)
    # the reference's CUDA kernel (StyleGAN3 filtered_lrelu.cu recompiled for sm_100a) on the same buffers: fp32 / fp16 only
    ref = ref_plugin('filtered_lrelu_plugin') if dtype in (torch.float32, torch.float16) else None
    if ref is not None:
        inf = float('inf')
        r_fwd = lambda i: ref.filtered_lrelu(xs[i % nrot], fu, fd, b, e, up, down, px0, px1, py0, py1, 0, 0, gain, slope, inf, False, True)
        r_bwd = lambda i: ref.filtered_lrelu(dys[i % nrot], fd, fu, zb, masks[i % nrot], down, up, pp[0], pp[1], pp[2], pp[3],
                                             px0 - (tu - 1), py0 - (tu - 1), gg, slope, inf, True, False)
        yr, sor, rcr = r_fwd(0)
        if rcr == 0:
            dxr = r_bwd(0)[0]
            dxo = bwd(0)[0]
            den = lambda t_: float(t_.float().abs().max().clamp_min(1e-30))
            rf.update(ref_cuda_us=round(time_graph(r_fwd, nrot, reps), 2), y_maxrel_vs_ref_cuda=float((y.float() - yr.float()).abs().max()) / den(yr),
                      mask_bytes_differ_vs_ref_cuda=int((sor != so).sum()), mask_bytes=int(so.numel()))
            rb.update(ref_cuda_us=round(time_graph(r_bwd, nrot, reps), 2), dx_maxrel_vs_ref_cuda=float((dxo.float() - dxr.float()).abs().max()) / den(dxr))
            for r_ in (rf, rb):
                r_['vs_ref_cuda'] = round(r_['ref_cuda_us'] / r_['us'], 3)
        else:
            rf['ref_cuda_us'] = rb['ref_cuda_us'] = None      # the reference has no specialised kernel for this call


def sweep_upfirdn2d(rows, N, C, S, mode, dtype, reps):
    f = cno_model.LReLu.design_lowpass_filter(12, S / 2.0001, 2 * (0.8 * S - S / 2.0001), 2 * S).to(DEV)
    kw = dict(up=2, padding=[6, 5, 6, 5], gain=4) if mode == 'up2' else dict(down=2, padding=[5, 5, 5, 5])
    es = torch.empty(0, dtype=dtype).element_size()
    so = S * 2 if mode == 'up2' else S // 2
    nbytes = N * C * (S * S + so * so) * es
    nrot = rot_count(nbytes)
    xs = [torch.randn(N, C, S, S, device=DEV, dtype=dtype) for _ in range(nrot)]
    fwd = lambda i: cno.upfirdn2d.upfirdn2d(xs[i % nrot], f, **kw)
    assert fwd(0).shape[-1] == so
    rows.append(row('upfirdn2d ' + mode, f'[{N},{C},{S}->{so}]', dtype, 'fwd', time_graph(fwd, nrot, reps), nbytes))


def sweep_bias_act(rows, N, C, S, dtype, reps):
    plug = cno.custom_ops.get_plugin('bias_act_plugin')
    es = torch.empty(0, dtype=dtype).element_size()
    n = N * C * S * S
    nrot = rot_count(3 * n * es)
    xs = [torch.randn(N, C, S, S, device=DEV, dtype=dtype) for _ in range(nrot)]
    dys = [torch.randn(N, C, S, S, device=DEV, dtype=dtype) for _ in range(nrot)]
    b = torch.randn(C, device=DEV, dtype=dtype); e = torch.empty(0)
    ys = [plug.bias_act(x, b, e, e, e, 0, 1, 3, 0.2, math.sqrt(2), -1) for x in xs]
    fwd = lambda i: plug.bias_act(xs[i % nrot], b, e, e, e, 0, 1, 3, 0.2, math.sqrt(2), -1)
    bwd = lambda i: plug.bias_act(dys[i % nrot], b, e, ys[i % nrot], e, 1, 1, 3, 0.2, math.sqrt(2), -1)
    shape = f'[{N},{C},{S},{S}] lrelu'
    rows.append(row('bias_act', shape, dtype, 'fwd', time_graph(fwd, nrot, reps), 2 * n * es))
    rows.append(row('bias_act', shape, dtype, 'bwd', time_graph(bwd, nrot, reps), 3 * n * es))


def sweep_cin(rows, N, C, S, dtype, reps):
    """FILM forward = cno_film_fwd (one kernel); backward = cno_cin_bwd + cno_film_bwd_finish, through the C ABI."""
    from pregenerating_pde_b200.torch_utils.custom_ops import _ptr, _stream, _DT
    lib = cno._lib.load()
    film = cno_model.FILM(C, dim=[2, 3], s=S).to(DEV)
    with torch.no_grad():
        for p_ in film.parameters():
            p_.add_(0.1 * torch.randn_like(p_))
    es = torch.empty(0, dtype=dtype).element_size()
    n = N * C * S * S
    nrot = rot_count(3 * n * es)
    xs = [torch.randn(N, C, S, S, device=DEV, dtype=dtype) for _ in range(nrot)]
    dys = [torch.randn(N, C, S, S, device=DEV, dtype=dtype) for _ in range(nrot)]
    ys = [torch.empty_like(x) for x in xs]
    t = torch.rand(N, device=DEV)
    P = [film.norm.weight, film.norm.bias, film.inp2lat_sacale.weight.reshape(-1), film.inp2lat_sacale.bias, film.lat2scale.weight,
         film.lat2scale.bias, film.inp2lat_bias.weight.reshape(-1), film.inp2lat_bias.bias, film.lat2bias.weight, film.lat2bias.bias]
    P = [v.detach().contiguous() for v in P]
    lat = P[2].numel()
    mean = torch.empty(N, C, device=DEV); rstd = torch.empty_like(mean); scale = torch.empty_like(mean)
    s12 = torch.empty(2, N, C, device=DEV)
    sizes = [C, C, lat, lat, C * lat, C, lat, lat, C * lat, C]
    g = list(torch.empty(sum(sizes), device=DEV).split(sizes))

    def fwd(i):
        rc = lib.cno_film_fwd(_ptr(xs[i % nrot]), _ptr(ys[i % nrot]), _ptr(t), *[_ptr(v) for v in P], _ptr(scale), _ptr(mean), _ptr(rstd),
                              _DT[dtype], N, C, S * S, lat, 1e-5, _stream(t))
        assert rc == 0

    def bwd(i):
        x = xs[i % nrot]
        rc = lib.cno_cin_bwd(_ptr(x), _ptr(dys[i % nrot]), _ptr(ys[i % nrot]), _ptr(P[0]), _ptr(scale), _ptr(mean), _ptr(rstd),
                             _ptr(s12[0]), _ptr(s12[1]), _DT[dtype], N, C, S * S, _stream(t))
        assert rc == 0
        rc = lib.cno_film_bwd_finish(_ptr(s12[0]), _ptr(s12[1]), _ptr(t), _ptr(P[0]), _ptr(P[1]), _ptr(scale), _ptr(P[2]), _ptr(P[3]),
                                     _ptr(P[4]), _ptr(P[6]), _ptr(P[7]), _ptr(P[8]), _ptr(g[0]), _ptr(g[1]), _ptr(g[4]), _ptr(g[5]),
                                     _ptr(g[2]), _ptr(g[3]), _ptr(g[8]), _ptr(g[9]), _ptr(g[6]), _ptr(g[7]), N, C, lat, _stream(t))
        assert rc == 0
    us_f = time_graph(fwd, nrot, reps)
    fwd(0)                               # statistics of buffer 0 for the backward rows (values do not affect timing)
    us_b = time_graph(bwd, nrot, reps)
    shape = f'[{N},{C},{S},{S}] FILM'
    rows.append(row('cin (FILM fwd, 1 kernel)', shape, dtype, 'fwd', us_f, 2 * n * es))
    rows.append(row('cin (FILM bwd, 2 kernels)', shape, dtype, 'bwd', us_b, 3 * n * es))


def sweep_film_lrelu(rows, N, C, S, dtype, reps):
    """LReLu(FILM(x, t)) forward, I-config, sign mask written: the prologue-fused pair (cno_film_stats + cno_filtered_lrelu_affine,
    no normalised tensor in HBM) beside the two-kernel form (cno_film_fwd + cno_filtered_lrelu).  Algorithmic bytes of both rows =
    x read once + y + mask written."""
    from pregenerating_pde_b200.torch_utils.custom_ops import _ptr, _stream, _DT
    lib = cno._lib.load()
    plug = cno.custom_ops.get_plugin('filtered_lrelu_plugin')
    film = cno_model.FILM(C, dim=[2, 3], s=S).to(DEV)
    act = cno_model.lrelu_for(C, S, S).to(DEV)
    up, down = act.up_factor, act.down_factor
    px0, px1, py0, py1 = act.padding
    fu, fd = act.up_filter, act.down_filter
    es = torch.empty(0, dtype=dtype).element_size()
    sh = S * down - (down - 1) + (fd.numel() - 1)
    swb = ((sh + 15) // 16 * 16) // 4
    nbytes = N * C * (2 * S * S * es + sh * swb) + C * es
    nrot = rot_count(nbytes + N * C * S * S * es)
    xs = [torch.randn(N, C, S, S, device=DEV, dtype=dtype) for _ in range(nrot)]
    zs = [torch.empty_like(x) for x in xs]
    t = torch.rand(N, device=DEV)
    P = [film.norm.weight, film.norm.bias, film.inp2lat_sacale.weight.reshape(-1), film.inp2lat_sacale.bias, film.lat2scale.weight,
         film.lat2scale.bias, film.inp2lat_bias.weight.reshape(-1), film.inp2lat_bias.bias, film.lat2bias.weight, film.lat2bias.bias]
    P = [v.detach().contiguous() for v in P]
    lat = P[2].numel()
    mean, rstd, scale, fa, fb = (torch.empty(N, C, device=DEV) for _ in range(5))
    b = torch.randn(C, device=DEV, dtype=dtype)
    e = torch.empty(0)
    gain, slope = math.sqrt(2), 0.2

    def fused(i):
        x = xs[i % nrot]
        rc = lib.cno_film_stats(_ptr(x), _ptr(t), *[_ptr(v) for v in P], _ptr(scale), _ptr(mean), _ptr(rstd), _ptr(fa), _ptr(fb),
                                _DT[dtype], N, C, S * S, lat, 1e-5, _stream(t))
        assert rc == 0
        assert plug.filtered_lrelu(x, fu, fd, b, e, up, down, px0, px1, py0, py1, 0, 0, gain, slope, float('inf'), False, True,
                                   affine=(fa, fb))[2] == 0

    def two(i):
        x, z = xs[i % nrot], zs[i % nrot]
        rc = lib.cno_film_fwd(_ptr(x), _ptr(z), _ptr(t), *[_ptr(v) for v in P], _ptr(scale), _ptr(mean), _ptr(rstd),
                              _DT[dtype], N, C, S * S, lat, 1e-5, _stream(t))
        assert rc == 0
        assert plug.filtered_lrelu(z, fu, fd, b, e, up, down, px0, px1, py0, py1, 0, 0, gain, slope, float('inf'), False, True)[2] == 0
    def stats(i):
        assert lib.cno_film_stats(_ptr(xs[i % nrot]), _ptr(t), *[_ptr(v) for v in P], _ptr(scale), _ptr(mean), _ptr(rstd), _ptr(fa), _ptr(fb),
                                  _DT[dtype], N, C, S * S, lat, 1e-5, _stream(t)) == 0

    def flr_aff(i):
        assert plug.filtered_lrelu(xs[i % nrot], fu, fd, b, e, up, down, px0, px1, py0, py1, 0, 0, gain, slope, float('inf'), False, True,
                                   affine=(fa, fb))[2] == 0
    shape = f'[{N},{C},{S},{S}] I'
    rows.append(row('film stats (read-only pass)', shape, dtype, 'fwd', time_graph(stats, nrot, reps), N * C * S * S * es))
    rows.append(row('filtered_lrelu, affine in', shape, dtype, 'fwd', time_graph(flr_aff, nrot, reps), nbytes))
    rows.append(row('film+lrelu fwd, fused pair', shape, dtype, 'fwd', time_graph(fused, nrot, reps), nbytes))
    rows.append(row('film+lrelu fwd, two ops', shape, dtype, 'fwd', time_graph(two, nrot, reps), nbytes))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--quick', action='store_true')
    ap.add_argument('--grid', action='store_true', help='also the full configs[1] grid: C in {16..256} x {128,64,32}^2 x {fp32,bf16}, I-config')
    ap.add_argument('--only', default='', help='comma list of op groups: flr,wide,up,bias,cin,pair (default: all)')
    ap.add_argument('--out', default=os.path.join(ROOT, 'profiles', 'sweep_latest.json'))
    ap.add_argument('--sweep', action='store_true')      # accepted when called through bench.py --sweep
    args, _ = ap.parse_known_args()
    cno._lib.load()
    N = 32
    rows = []
    levels = [(16, 128), (32, 64), (64, 32), (128, 16), (256, 8)]          # (C, H) of the CNO U-Net levels
    dtypes = [torch.float32] if args.quick else [torch.float32, torch.bfloat16]
    only = set(filter(None, args.only.split(',')))
    want = lambda g: not only or g in only
    for dt in dtypes:
        reps = 10
        if want('flr'):
            sweep_filtered_lrelu(rows, N, 64, 128, 128, dt, reps)             # the bench.py workload (lift / project shape)
            for C, H in levels:
                sweep_filtered_lrelu(rows, N, C, H, H, dt, reps)              # I
            for (C, H) in [(32, 128), (64, 64), (128, 32), (256, 16)]:
                sweep_filtered_lrelu(rows, N, C, H, H // 2, dt, reps)         # D
            for (C, H) in [(128, 8), (64, 16), (32, 32), (16, 64)]:
                sweep_filtered_lrelu(rows, N, C, H, H * 2, dt, reps)          # U
        if want('wide'):                                                      # BASELINE.json configs[4]: 256^2 fine-tune planes
            sweep_filtered_lrelu(rows, 8, 64, 256, 256, dt, reps)
            sweep_filtered_lrelu(rows, 8, 64, 256, 128, dt, reps)
            sweep_filtered_lrelu(rows, 8, 32, 128, 256, dt, reps)
        if args.grid:                                                         # BASELINE.json configs[1]: the whole C x size grid
            for H in (128, 64, 32):
                for C in (16, 32, 64, 128, 256):
                    if not any(r['shape'] == f'[{N},{C},{H}->{H}] I' and r['dtype'] == str(dt).replace('torch.', '') for r in rows):
                        sweep_filtered_lrelu(rows, N, C, H, H, dt, reps)
        if want('up'):
            for C, H in [(64, 128), (16, 128), (64, 32)]:
                sweep_upfirdn2d(rows, N, C, H, 'up2', dt, reps)
                sweep_upfirdn2d(rows, N, C, H, 'down2', dt, reps)
        if want('bias'):
            for C, H in [(64, 128), (64, 32)]:
                sweep_bias_act(rows, N, C, H, dt, reps)
        if want('cin'):
            for C, H in levels + [(64, 128)]:
                sweep_cin(rows, N, C, H, dt, reps)
            sweep_cin(rows, 8, 64, 256, dt, reps)                             # wide fine-tune planes: thread-block cluster kernels
        if want('pair'):
            for C, H in [(64, 128), (16, 128), (64, 32)]:
                sweep_film_lrelu(rows, N, C, H, dt, reps)                     # CIN -> filtered_lrelu prologue fusion
    out = dict(gpu=torch.cuda.get_device_name(0), hbm_peak_gbs=peak(), rows=rows)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(out, open(args.out, 'w'), indent=1)
    print(f'{"op":28s} {"shape":26s} {"dtype":9s} {"dir":4s} {"us":>9s} {"GB/s":>8s} {"frac":>6s} {"ref_cuda_us":>12s} {"x":>6s}')
    for r in rows:
        extra = f' {r["ref_cuda_us"]:12.2f} {r["vs_ref_cuda"]:6.2f}' if r.get('ref_cuda_us') else ''
        print(f'{r["op"]:28s} {r["shape"]:26s} {r["dtype"]:9s} {r["dir"]:4s} {r["us"]:9.2f} {r["gbs"]:8.1f} {r["frac_hbm"]:6.3f}{extra}')


if __name__ == '__main__':
    main()
