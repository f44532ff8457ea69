#!/usr/bin/env python
"""bench.py -- headline measurement of the CNO activation hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W]            # our arm (one JSON line on stdout)
    python bench.py --impl reference [--steps K] [--warmup W]      # the reference's CPU impl='ref' algorithm
    python bench.py --sweep                                         # per-op table -> profiles/sweep_latest.json

Step = one filtered_lrelu forward + backward (dx, db) over one synthetic batch of planes
(BASELINE.json configs[1], row fp32 [32,64,128,128], CNO 'I' configuration: up 2 / 12 taps, down 2 /
12 taps, gain sqrt2, slope 0.2).  `value` = algorithmic GB/s with inputs resident in HBM;
`e2e` = the same metric through the public API with HOST (pinned) buffers, copies inside the timed
region.  Multi-GPU: planes are sharded by batch, no data-path collective (weak scaling).
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

WORKLOAD = dict(name='filtered_lrelu fwd+bwd fp32 [32,64,128,128] I-config (up2/12 taps, down2/12 taps)',
                N=32, C=64, S_in=128, S_out=128)
CPU_SAMPLE = dict(N=8, C=64, S_in=128, S_out=128)   # BASELINE.json configs[0]: the reference's CPU-runnable case


def algorithmic_bytes(N, C, H, OH, e, up_taps_d, down):
    """fwd = N*C*(H*W*e + OH*OW*e + sh*swb) + C*e ; bwd the same (+ C*e for db)."""
    sh = OH * down - (down - 1) + (up_taps_d - 1)
    swb = ((OH * down - (down - 1) + (up_taps_d - 1) + 15) // 16 * 16) // 4
    per_plane = H * H * e + OH * OH * e + sh * swb
    return N * C * per_plane + C * e, N * C * per_plane + C * e


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs'], 'measured (MEASURED_PEAKS.json)'
    except Exception:
        return 6650.0, 'fallback (B200_PROFILING.md)'


class ClockSampler:
    Q = 'clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,' \
        'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap'

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', f'--id={self.index}', f'--query-gpu={self.Q}', '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True); self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def stop(self):
        if not self.proc:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=['nvidia-smi unavailable'])
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for r in self.rows if len(r) >= 6 and r[0].isdigit()]
        if not rows:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=['no samples'])
        sm = sorted(int(r[0]) for r in rows)
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith('active') for r in rows)]
        return dict(sm_mhz=sm[len(sm) // 2], sm_max_mhz=int(rows[0][1]), reasons=reasons, samples=len(rows))


def oracle_step(O, cfg, x, b, dy):
    xr, br = x.requires_grad_(True), b.requires_grad_(True)
    y = O.filtered_lrelu(xr, cfg['fu'], cfg['fd'], br, up=cfg['up'], down=cfg['down'], padding=cfg['padding'])
    return torch.autograd.grad(y, [xr, br], dy)


def time_cpu(steps, warmup, sample):
    """the reference algorithm (oracle port of impl='ref', torch conv2d based) on the host cores"""
    from oracle import cno_oracle as O
    torch.set_num_threads(os.cpu_count())
    cfg = O.lrelu_config(sample['S_in'], sample['S_out'])
    g = torch.Generator().manual_seed(0)
    x = torch.randn(sample['N'], sample['C'], sample['S_in'], sample['S_in'], generator=g)
    b = torch.randn(sample['C'], generator=g)
    dy = torch.randn(sample['N'], sample['C'], sample['S_out'], sample['S_out'], generator=g)
    for _ in range(warmup):
        oracle_step(O, cfg, x.detach(), b.detach(), dy)
    t0 = time.perf_counter()
    for _ in range(steps):
        oracle_step(O, cfg, x.detach(), b.detach(), dy)
    dt = (time.perf_counter() - t0) / steps
    fb, bb = algorithmic_bytes(sample['N'], sample['C'], sample['S_in'], sample['S_out'], 4, len(cfg['fd']), cfg['down'])
    return (fb + bb) / dt / 1e9, dt, torch.get_num_threads()


def run_reference_arm(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    steps, warmup = max(1, min(args.steps, 10)), max(1, min(args.warmup, 2))
    gbs, dt, cores = time_cpu(steps, warmup, CPU_SAMPLE)
    sample = f"oracle port of impl='ref' on fp32 [{CPU_SAMPLE['N']},{CPU_SAMPLE['C']},128,128] planes (1/4 of the batch), {steps} timed passes"
    print(json.dumps({
        'impl': 'reference', 'metric': 'filtered_lrelu_fwd_bwd_algorithmic_GBps', 'value': round(gbs, 4), 'unit': 'GB/s',
        'n_gpus': args.gpus, 'steps': steps, 'warmup': warmup, 'ms_per_step': round(dt * 1e3, 3), 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': WORKLOAD['name'], 'sample': sample},
        'cpu_baseline': {'value': round(gbs, 4), 'unit': 'GB/s', 'cores': cores, 'kind': 'port', 'sample': sample},
        'e2e': {'value': round(gbs, 4), 'unit': 'GB/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    }))


def _finish(world, dev):
    """Multi-rank exit.  Tearing down an NCCL communicator whose collectives live in captured CUDA graphs can block in the
    communicator destructor (seen at 8 ranks), so after a last barrier every rank leaves without running destructors."""
    if world > 1:
        import torch.distributed as dist
        torch.cuda.synchronize(dev)
        dist.barrier()
        torch.cuda.synchronize(dev)
        sys.stdout.flush(); sys.stderr.flush()
        os._exit(0)


def run_ours(args):
    import torch.distributed as dist
    import pregenerating_pde_b200 as cno
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    cno._lib.load()
    W = WORKLOAD
    act = cno.cno_model.lrelu_for(W['C'], W['S_in'], W['S_out']).to(dev)      # filters / padding exactly as CNOBlock builds them
    cfg = dict(up=act.up_factor, down=act.down_factor, padding=act.padding, fd=act.down_filter)
    fu, fd = act.up_filter, act.down_filter
    g = torch.Generator(device=dev).manual_seed(rank)
    x = torch.randn(W['N'], W['C'], W['S_in'], W['S_in'], device=dev, generator=g)
    b = torch.randn(W['C'], device=dev, generator=g)
    dy = torch.randn(W['N'], W['C'], W['S_out'], W['S_out'], device=dev, generator=g)
    F = cno.filtered_lrelu.filtered_lrelu
    kw = dict(up=cfg['up'], down=cfg['down'], padding=cfg['padding'])

    def step(xd, bd, dyd):
        xd.requires_grad_(True); bd.requires_grad_(True)
        y = F(xd, fu, fd, bd, **kw)
        return torch.autograd.grad(y, [xd, bd], dyd)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    fb, bb = algorithmic_bytes(W['N'], W['C'], W['S_in'], W['S_out'], 4, len(cfg['fd']), cfg['down'])
    # two input sets used alternately; one step moves 611 MB through HBM, far more than the 126 MB L2
    pool = [(x, dy), (torch.randn_like(x), torch.randn_like(dy))]
    for i in range(max(args.warmup, 3)):
        step(pool[i % 2][0].detach(), b.detach(), pool[i % 2][1])
    # ---- kernel-only: device-resident inputs.  The step (public API: forward + autograd backward) is captured once per
    # input set into a CUDA graph and replayed, so the timed region holds GPU work only (no Python / ctypes launch gaps).
    def capture(fn):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            out = fn()
        return g, out
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        g_step = [capture(lambda i=i: step(pool[i][0].detach(), b.detach(), pool[i][1])) for i in range(2)]
        g_fwd = [capture(lambda i=i: F(pool[i][0].detach().requires_grad_(True), fu, fd, b, **kw)) for i in range(2)]
    torch.cuda.current_stream().wait_stream(side)
    for i in range(4):
        g_step[i % 2][0].replay(); g_fwd[i % 2][0].replay()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e_begin, e_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e_begin.record()
    for k in range(args.steps):
        g_step[k % 2][0].replay()
    e_end.record()
    barrier()
    ms_total = e_begin.elapsed_time(e_end)
    # the forward kernel alone (same graphs minus the backward), for the per-kernel launch durations of the roofline
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    for k in range(args.steps):
        g_fwd[k % 2][0].replay()
    f1.record()
    barrier()
    fwd_ms = f0.elapsed_time(f1) / args.steps
    bwd_ms = ms_total / args.steps - fwd_ms
    # ---- end to end: host (pinned) buffers through the public API, H2D + D2H inside the timed region.  The batch is cut
    # into chunks along N, each on its own stream, so the PCIe copies of one chunk overlap the kernels of another.
    hx, hdy = x.cpu().pin_memory(), dy.cpu().pin_memory()
    hb = b.cpu().pin_memory()
    hdx, hdb = torch.empty_like(hx).pin_memory(), torch.empty_like(hb).pin_memory()
    e2e_steps = max(2, min(args.steps, 10))
    NCH = 8
    streams = [torch.cuda.Stream() for _ in range(NCH)]
    cn = W['N'] // NCH
    dbs = torch.zeros(NCH, W['C'], device=dev)

    def e2e_step():
        cur = torch.cuda.current_stream()
        bd = hb.to(dev, non_blocking=True)
        for c, s_ in enumerate(streams):
            s_.wait_stream(cur)
            with torch.cuda.stream(s_):
                sl = slice(c * cn, (c + 1) * cn)
                xd = hx[sl].to(dev, non_blocking=True); dyd = hdy[sl].to(dev, non_blocking=True)
                dx, db = step(xd, bd.detach(), dyd)
                hdx[sl].copy_(dx, non_blocking=True)
                dbs[c].copy_(db)
        for s_ in streams:
            cur.wait_stream(s_)
        hdb.copy_(dbs.sum(0), non_blocking=True)
    e2e_step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(e2e_steps):
        e2e_step()
    e1.record()
    barrier()
    e2e_ms = e0.elapsed_time(e1) / e2e_steps
    clocks = sampler.stop() if rank == 0 else None
    # ---- max over ranks ----------------------------------------------------------------------------
    t = torch.tensor([ms_total, e2e_ms, fwd_ms, bwd_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, e2e_ms, fwd_ms, bwd_ms = t.tolist()
    if rank == 0:
        ms_step = ms_total / args.steps
        peak, peak_src = peaks()
        # per-kernel roofline: the step is two launches of the same fused kernel (forward = SIGN_WRITE, backward = SIGN_READ
        # with db accumulated in the kernel); the slower one is the dominant kernel
        try:
            ncu = json.load(open(os.path.join(ROOT, 'profiles', 'ncu_traffic.json')))
        except Exception:
            ncu = {}
        kern = {'fwd': dict(name='flr_fused_kernel<float, Cfg<2,2,1,1>, SIGN_WRITE> (forward, writes the sign mask)', ms=fwd_ms, bytes=fb,
                            traffic=ncu.get('flr_fwd_dram_bytes_per_launch')),
                'bwd': dict(name='flr_fused_kernel<float, Cfg<2,2,1,1>, SIGN_READ> (backward, reads the sign mask, accumulates db)', ms=bwd_ms,
                            bytes=bb, traffic=ncu.get('flr_bwd_dram_bytes_per_launch'))}
        dom = 'bwd' if bwd_ms >= fwd_ms else 'fwd'
        oth = 'fwd' if dom == 'bwd' else 'bwd'
        gbs = lambda k: kern[k]['bytes'] / (kern[k]['ms'] * 1e-3) / 1e9
        achieved = gbs(dom)
        traffic = kern[dom]['traffic']
        cpu_gbs, cpu_dt, cores = time_cpu(8, 1, CPU_SAMPLE)
        h2d = hx.numel() * 4 + hdy.numel() * 4 + hb.numel() * 4
        d2h = hdx.numel() * 4 + hdb.numel() * 4
        out = {
            'metric': 'filtered_lrelu_fwd_bwd_algorithmic_GBps', 'value': round(world * (fb + bb) / (ms_step * 1e-3) / 1e9, 2),
            'unit': 'GB/s', 'n_gpus': world, 'steps': args.steps, 'warmup': max(args.warmup, 3), 'ms_per_step': round(ms_step, 4),
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': {'workload': W['name'], 'planes_per_gpu': W['N'] * W['C'], 'algorithmic_bytes_per_step_per_gpu': fb + bb,
                       'l2': 'inputs larger than L2 (x + dy + mask = 305 MB per step > 126 MB), no explicit flush',
                       'fwd_ms': round(fwd_ms, 4), 'bwd_ms_incl_db': round(bwd_ms, 4), 'timed_as': 'CUDA graph replays of the public-API step',
                       'e2e_pipeline': f'{NCH} chunks along N on {NCH} streams (H2D / kernels / D2H overlap)'},
            'e2e': {'value': round(world * (fb + bb) / (e2e_ms * 1e-3) / 1e9, 2), 'unit': 'GB/s', 'h2d_bytes_per_step': h2d,
                    'd2h_bytes_per_step': d2h, 'ms_per_step': round(e2e_ms, 3)},
            'gpu_launches': 2 * args.steps,
            'roofline': {'kernel': kern[dom]['name'], 'bound': 'hbm',
                         'achieved': round(achieved, 1), 'peak': peak, 'unit': 'GB/s', 'frac': round(achieved / peak, 4),
                         'traffic': traffic, 'peak_source': peak_src, 'algorithmic_bytes_per_launch': kern[dom]['bytes'],
                         'avg_launch_ms': round(kern[dom]['ms'], 4),
                         'timing': 'CUDA events around K graph replays of the step and of the forward alone; backward = difference',
                         'other_kernel': {'kernel': kern[oth]['name'], 'achieved': round(gbs(oth), 1), 'frac': round(gbs(oth) / peak, 4),
                                          'avg_launch_ms': round(kern[oth]['ms'], 4), 'traffic': kern[oth]['traffic']},
                         'fp32_pipe_note': 'op is 16.5 FLOP/B: the FP32 FMA pipe bounds this launch at 75.5 us = 0.62 of the HBM roofline '
                                           '(DESIGN.md 3.1); ncu pipe_fma_cycles_active is reported in profiles/'},
            'cpu_baseline': {'value': round(cpu_gbs, 4), 'unit': 'GB/s', 'cores': cores, 'kind': 'port',
                             'sample': f"oracle port of impl='ref', fp32 [8,64,128,128] (1/4 of the batch), 8 timed passes, {cpu_dt * 1e3:.0f} ms each"},
            'clocks': clocks,
        }
        print(json.dumps(out), flush=True)
    _finish(world, dev)


def run_cno_train(args):
    """Secondary workload (BASELINE.json configs[2], [3]): CNO_time (TrainCNO_time_L configuration, 18.28 M parameters) training
    step on synthetic FPO batches, batch 32 per GPU, DDP over NCCL for N > 1; value = samples/s over all GPUs."""
    import torch.distributed as dist
    import pregenerating_pde_b200 as cno
    world = int(os.environ.get('WORLD_SIZE', '1')); rank = int(os.environ.get('RANK', '0')); local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    torch.backends.cudnn.benchmark = True
    B = args.batch
    tr = cno.train.Trainer(device=dev, ddp=world > 1, graph=bool(args.graph))
    batches = [cno.train.synthetic_fpo_batch(B, device='cpu', seed=100 * rank + i) for i in range(2)]
    host = [tuple(v.pin_memory() for v in b) for b in batches]
    devb = [tuple(v.to(dev) for v in b) for b in batches]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    for i in range(max(args.warmup, 3)):
        tr.step(*devb[i % 2])
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier(); e0.record()
    for i in range(args.steps):
        loss = tr.step(*devb[i % 2])
    e1.record(); barrier()
    ms = e0.elapsed_time(e1) / args.steps
    # end to end: batch from pinned host memory every step, loss read back every step
    barrier(); t0 = time.perf_counter()
    for i in range(args.steps):
        b = tuple(v.to(dev, non_blocking=True) for v in host[i % 2])
        lv = float(tr.step(*b))
    barrier(); e2e_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    clocks = sampler.stop() if rank == 0 else None
    tt = torch.tensor([ms, e2e_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms, e2e_ms = tt.tolist()
    if rank == 0:
        h2d = sum(v.numel() * v.element_size() for v in host[0])
        print(json.dumps({
            'metric': 'cno_train_samples_per_s', 'value': round(world * B / (ms * 1e-3), 2), 'unit': 'samples/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': max(args.warmup, 3), 'ms_per_step': round(ms, 3), 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': {'workload': 'CNO_time TrainCNO_time_L config (N_layers 4, N_res 8, 18.28M params) fwd + rel-L1 + bwd + AdamW, FPO 7ch 128x128',
                       'batch_per_gpu': B, 'cuda_graph': bool(tr.use_graph), 'parallelism': f'dp{world}', 'grad_allreduce': tr.dp_mode, 'final_loss': round(float(loss), 5)},
            'e2e': {'value': round(world * B / (e2e_ms * 1e-3), 2), 'unit': 'samples/s', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': 4,
                    'ms_per_step': round(e2e_ms, 3)},
            'clocks': clocks}), flush=True)
    _finish(world, dev)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--workload', default='flr', choices=['flr', 'cno_train'])
    ap.add_argument('--batch', type=int, default=32)
    ap.add_argument('--graph', type=int, default=1)
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=30)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--sweep', action='store_true')
    args = ap.parse_args()
    if args.sweep:
        import bench_sweep
        return bench_sweep.main()
    if args.impl == 'reference':
        return run_reference_arm(args)
    if args.workload == 'cno_train':
        return run_cno_train(args)
    return run_ours(args)


if __name__ == '__main__':
    main()
