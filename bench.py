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
    """Multi-rank exit.  Round 1 left with os._exit because destroying the NCCL communicator blocked at 8 ranks: the captured
    CUDA graphs that hold its kernels were still alive.  Now every graph is dropped first (Trainer.close / del), then the
    process group is destroyed in order.  A watchdog thread still ends the process if teardown ever blocks, so a hang cannot
    cost the driver its scaling run."""
    if world > 1:
        import torch.distributed as dist
        torch.cuda.synchronize(dev)
        dist.barrier()
        torch.cuda.synchronize(dev)
        sys.stdout.flush(); sys.stderr.flush()
        t = threading.Timer(20.0, lambda: os._exit(0))
        t.daemon = True
        t.start()
        dist.destroy_process_group()
        t.cancel()


def run_ours(args):
    import torch.distributed as dist
    import pregenerating_pde_b200 as cno
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    # CPU baseline first, on rank 0, while the other ranks sleep in the rendezvous below (a rank that spins in an NCCL
    # barrier next to it took half the host cores away: 0.53 -> 0.28 GB/s at N >= 2 in round 1)
    cpu_rec = time_cpu(8, 1, CPU_SAMPLE) if rank == 0 else None
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
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        sm_max = (clocks or {}).get('sm_max_mhz') or 1965
        fp32_floor_ms = W['N'] * W['C'] * 1234008 / (sms * 128 * sm_max * 1e6) * 1e3
        oth = 'fwd' if dom == 'bwd' else 'bwd'
        gbs = lambda k: kern[k]['bytes'] / (kern[k]['ms'] * 1e-3) / 1e9
        achieved = gbs(dom)
        traffic = kern[dom]['traffic']
        cpu_gbs, cpu_dt, cores = cpu_rec
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
                         'traffic_source': 'profiles/ncu_traffic.json (dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full capture of this kernel; not re-measured in this run)',
                         # the op is 16.5 FLOP/B: at fp32 the FMA pipe, not HBM, is the binding roof.  MACs per launch =
                         # planes x 1,234,008; pipe peak = SMs x 128 lanes x max SM clock
                         'fp32_pipe_floor_ms': round(fp32_floor_ms, 4), 'fp32_pipe_frac': round(fp32_floor_ms / kern[dom]['ms'], 4),
                         'compound_frac': round(max(fp32_floor_ms, kern[dom]['bytes'] / peak / 1e6) / kern[dom]['ms'], 4),
                         'compound_note': 'time / max(HBM time, FP32-FMA time): the ceiling of frac for a perfect CUDA-core kernel is '
                                          f"{round(kern[dom]['bytes'] / peak / 1e6 / fp32_floor_ms, 3)}"},
            'cpu_baseline': {'value': round(cpu_gbs, 4), 'unit': 'GB/s', 'cores': cores, 'kind': 'port',
                             'sample': f"oracle port of impl='ref', fp32 [8,64,128,128] (1/4 of the batch), 8 timed passes, {cpu_dt * 1e3:.0f} ms each"},
            'clocks': clocks,
        }
    # ---- BASELINE.json's second metric in the same line: CNO_time training samples/s (gradient all-reduce over NCCL at N > 1)
    del g_step, g_fwd, pool, x, dy, hx, hdy, hdx
    torch.cuda.empty_cache()
    train_rec = None
    if not args.no_train:
        targs = argparse.Namespace(**vars(args)); targs.steps = max(3, min(args.steps, 20))
        train_rec = measure_train(targs, world, rank, dev, mode='train')
    if rank == 0:
        out['cno_train'] = train_rec
        print(json.dumps(out), flush=True)
    _finish(world, dev)


def measure_train(args, world, rank, dev, mode='train'):
    """CNO_time training step (TrainCNO_time_L configuration) or the wide fine-tune step (BASELINE.json configs[4]) on synthetic
    batches, batch `args.batch` per GPU; returns the record rank 0 prints (None on the other ranks).  The process group is
    already initialised by the caller when world > 1."""
    import torch.distributed as dist
    import pregenerating_pde_b200 as cno
    torch.backends.cudnn.benchmark = True
    B = args.batch
    if mode == 'finetune':
        ft = cno.finetune.wide_finetune_config()
        tr = cno.train.Trainer(device=dev, ddp=world > 1, graph=bool(args.graph), mode='finetune', ft=ft,
                               autocast=torch.bfloat16 if args.autocast == 'bf16' else None)
        batches = [cno.train.synthetic_finetune_batch(B, ft, device='cpu', seed=100 * rank + i) for i in range(2)]
        name = ('CNO_FineTune wide config: foundation CNO_time (in 5 / out 4, N_layers 4, N_res 8) at 256x256, channel_multiplier 64 '
                '(features 32..512), initialize_FT(5->4 in, 4->3 out: 128x128 output), 3-group AdamW')
    else:
        tr = cno.train.Trainer(device=dev, ddp=world > 1, graph=bool(args.graph),
                               autocast=torch.bfloat16 if args.autocast == 'bf16' else None)
        batches = [cno.train.synthetic_fpo_batch(B, device='cpu', seed=100 * rank + i) for i in range(2)]
        name = 'CNO_time TrainCNO_time_L config (N_layers 4, N_res 8, 18.28M params) fwd + rel-L1 + bwd + AdamW, FPO 7ch 128x128'
    nparam = sum(p.numel() for p in tr.params)
    host = [tuple(v.pin_memory() for v in b) for b in batches]
    devb = [tuple(v.to(dev) for v in b) for b in batches]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    for i in range(max(args.warmup, 3)):
        tr.step(*devb[i % 2])
    sampler = ClockSampler(dev.index or 0)
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
    rec = None
    if rank == 0:
        h2d = sum(v.numel() * v.element_size() for v in host[0])
        rec = {'metric': 'cno_train_samples_per_s' if mode == 'train' else 'cno_finetune_samples_per_s',
               'value': round(world * B / (ms * 1e-3), 2), 'unit': 'samples/s', 'n_gpus': world,
               'steps': args.steps, 'warmup': max(args.warmup, 3), 'ms_per_step': round(ms, 3), 'higher_is_better': True, 'scaling': 'weak',
               'vs_baseline': None, 'dtype': 'bf16 autocast (convolutions), fp32 master weights' if args.autocast == 'bf16' else 'f32', 'data': 'synthetic',
               'config': {'workload': name, 'parameters': nparam, 'batch_per_gpu': B, 'cuda_graph': bool(tr.use_graph), 'parallelism': f'dp{world}',
                          'grad_allreduce': tr.dp_mode, 'grad_bytes': nparam * 4, 'final_loss': round(float(loss), 5)},
               'e2e': {'value': round(world * B / (e2e_ms * 1e-3), 2), 'unit': 'samples/s', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': 4,
                       'ms_per_step': round(e2e_ms, 3)},
               'clocks': clocks}
    tr.close()
    del tr
    torch.cuda.empty_cache()
    return rec


def run_cno_rollout(args):
    """Forward-only autoregressive rollout of TestCNO_ALL.py:124-188 (`pred = model(inp, t)`; `inp[:, :dim] = pred`), K = 7 steps of
    the TrainCNO_time_L model captured as ONE CUDA graph under no_grad (SIGN_NONE launches: no sign masks are written);
    value = model evaluations per second (samples x rollout steps)."""
    import pregenerating_pde_b200 as cno
    torch.cuda.set_device(0)
    dev = torch.device('cuda', 0)
    torch.backends.cudnn.benchmark = True
    model = cno.cno_model.CNO_time(**cno.cno_model.train_config_L()).to(dev).eval()
    B, K = args.batch, 7
    t, x, _ = cno.train.synthetic_fpo_batch(B, device=dev, seed=1)
    hx = x.cpu().pin_memory()

    def rollout(inp):
        outs = []
        for _ in range(K):
            pred = model(inp, t)
            inp = torch.cat([pred, inp[:, pred.shape[1]:]], 1)
            outs.append(pred)
        return torch.stack(outs)
    with torch.no_grad():
        side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                rollout(x)
        torch.cuda.current_stream().wait_stream(side)
        static_x = x.clone()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            out = rollout(static_x)
        for _ in range(max(args.warmup, 3)):
            g.replay()
        sampler = ClockSampler(0); sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(args.steps):
            g.replay()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
        t0 = time.perf_counter()
        for _ in range(args.steps):                       # e2e: initial condition from pinned host memory, last prediction's mean read back
            static_x.copy_(hx, non_blocking=True); g.replay(); v = float(out[-1].mean())
        e2e_ms = (time.perf_counter() - t0) * 1e3 / args.steps
        clocks = sampler.stop()
    print(json.dumps({'metric': 'cno_rollout_model_evaluations_per_s', 'value': round(B * K / (ms * 1e-3), 1), 'unit': 'samples*steps/s', 'n_gpus': 1,
                      'steps': args.steps, 'warmup': max(args.warmup, 3), 'ms_per_step': round(ms, 3), 'higher_is_better': True, 'scaling': 'weak',
                      'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
                      'config': {'workload': f'CNO_time TrainCNO_time_L model, no_grad autoregressive rollout of {K} steps (TestCNO_ALL.py:124-188), batch {B}, one CUDA graph',
                                 'rollout_steps': K, 'batch': B},
                      'e2e': {'value': round(B * K / (e2e_ms * 1e-3), 1), 'unit': 'samples*steps/s', 'h2d_bytes_per_step': hx.numel() * 4, 'd2h_bytes_per_step': 4,
                              'ms_per_step': round(e2e_ms, 3)}, 'clocks': clocks}), flush=True)


def run_cno_train(args, mode='train'):
    """Secondary workloads: `--workload cno_train` (BASELINE.json configs[2], [3]) and `--workload cno_finetune` (configs[4]) as
    the headline of their own JSON line; value = samples/s over all GPUs."""
    import torch.distributed as dist
    world = int(os.environ.get('WORLD_SIZE', '1')); rank = int(os.environ.get('RANK', '0')); local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    rec = measure_train(args, world, rank, dev, mode=mode)
    if rank == 0:
        print(json.dumps(rec), flush=True)
    _finish(world, dev)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--workload', default='flr', choices=['flr', 'cno_train', 'cno_finetune', 'cno_rollout'])
    ap.add_argument('--autocast', default='none', choices=['none', 'bf16'])
    ap.add_argument('--no-train', dest='no_train', action='store_true', help='flr workload: skip the cno_train sub-record')
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
    if args.workload == 'cno_rollout':
        return run_cno_rollout(args)
    if args.workload == 'cno_finetune':
        if args.batch == 32 and args.autocast == 'none':
            args.autocast = 'bf16'                     # BASELINE.json configs[4]: bf16
        return run_cno_train(args, mode='finetune')
    return run_ours(args)


if __name__ == '__main__':
    main()
