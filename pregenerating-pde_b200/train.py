"""CNO training step harness: the workload of TrainCNO_time_L.py (CNO_Experiments/TrainCNO_time_L.py:167-224 builds
the model and hands it to a Lightning Trainer with DDP) restated without Lightning: synthetic FPO-shaped batches,
forward + relative-L1 loss + backward + AdamW, one process per GPU with torch DDP over NCCL (gradient all-reduce is
the only exchange), optional whole-step CUDA-graph capture on a single GPU.
"""
import math
import contextlib
import os
import torch
import torch.distributed as dist
from torch import nn

from . import cno_model


def synthetic_fpo_batch(batch, size=128, device='cuda', seed=0, in_dim=7, out_dim=3, dtype=torch.float32):
    """FPO-shaped sample layout of NSFlowTimeDataset.__getitem__ (DataLoaders/CNO_TimeLoaders.py:219-248):
    x[B,7,S,S] = (Ux, Uy, p z-scored | Re plane | mask | SDF | lead-time plane), t[B] = k/19, y[B,3,S,S] z-scored."""
    g = torch.Generator(device='cpu').manual_seed(seed)
    S = size
    uvp = torch.stack([1.0 + 0.5 * torch.randn(batch, S, S, generator=g), 0.1 + 0.3 * torch.randn(batch, S, S, generator=g),
                       10.0 + 2.0 * torch.randn(batch, S, S, generator=g)], 1)
    mean, std = uvp.mean((0, 2, 3), keepdim=True), uvp.std((0, 2, 3), keepdim=True)
    uvp = (uvp - mean) / std
    yy, xx = torch.meshgrid(torch.arange(S, dtype=torch.float32), torch.arange(S, dtype=torch.float32), indexing='ij')
    r = torch.sqrt((xx - S / 2) ** 2 + (yy - S / 2) ** 2)
    mask = (r < S / 6).float()
    sdf = ((r - S / 6) / (S / 2)).clamp(-1, 1)
    re = torch.rand(batch, generator=g)
    k = torch.randint(1, 20, (batch,), generator=g).float()
    t = k / 19.0
    planes = [uvp, re.view(-1, 1, 1, 1).expand(-1, 1, S, S), mask.expand(batch, 1, S, S), sdf.expand(batch, 1, S, S),
              t.view(-1, 1, 1, 1).expand(-1, 1, S, S)]
    x = torch.cat(planes, 1)[:, :in_dim] if in_dim <= 7 else None
    tgt = (uvp + 0.1 * torch.randn(batch, 3, S, S, generator=g))[:, :out_dim]
    return t.to(device), x.contiguous().to(device, dtype), tgt.contiguous().to(device, dtype)


class DeviceNSFlowBatcher:
    """Device-resident sample construction of NSFlowTimeDataset (DataLoaders/CNO_TimeLoaders.py:183-248): the whole
    trajectory array [N, T, H, W, 6] lives in HBM (the FPO set is a few GB; a B200 has 180 GB) and a batch of dataset indices
    becomes (time, input, label) with three gathers and two fused elementwise ops on the GPU instead of a per-sample
    `__getitem__` on the host plus a collate and an H2D copy.  Index arithmetic, z-scoring of (ux, uy, p) only, the lead-time
    plane appended last and time = (t2 - t1) / 19 are the reference's."""

    def __init__(self, data, time_indices, start=0, mean=None, std=None, time_input=True, device='cuda'):
        self.data = torch.as_tensor(data, dtype=torch.float32).to(device)          # [N, T, H, W, 6]
        self.pairs = torch.as_tensor(time_indices, dtype=torch.long, device=device)   # [multiplier, 2] = (t1, t2)
        self.multiplier, self.start, self.time_input = len(time_indices), int(start), bool(time_input)
        if mean is None or std is None:                                             # per-file statistics (:205-208)
            mean = self.data[..., :3].mean(dim=(0, 1, 2, 3)); std = self.data[..., :3].std(dim=(0, 1, 2, 3), unbiased=False)
        self.mean = torch.as_tensor(mean, dtype=torch.float32, device=device).view(1, 3, 1, 1)
        self.std = torch.as_tensor(std, dtype=torch.float32, device=device).view(1, 3, 1, 1)

    def batch(self, idx):
        idx = torch.as_tensor(idx, dtype=torch.long, device=self.data.device)
        i, k = idx // self.multiplier + self.start, idx % self.multiplier
        t1, t2 = self.pairs[k, 0], self.pairs[k, 1]
        time = (t2 - t1).float() / 19.0
        inp = self.data[i, t1].permute(0, 3, 1, 2)                                  # [B, 6, H, W]
        lab = self.data[i, t2][..., :3].permute(0, 3, 1, 2)
        inp = torch.cat([(inp[:, :3] - self.mean) / self.std, inp[:, 3:]], 1)
        lab = (lab - self.mean) / self.std
        if self.time_input:
            inp = torch.cat([inp, time.view(-1, 1, 1, 1).expand(-1, 1, inp.shape[2], inp.shape[3])], 1)
        return time, inp.contiguous(), lab.contiguous()


def freeze_dead_parameters(model):
    """Parameters that never receive a gradient are marked frozen: AdamW skips a parameter without gradient either way, so this
    changes nothing numerically, but DDP needs no unused-parameter search and the gradient buckets complete during backward.
    (1) decoder_inv[N_layers] is constructed but never called (CNO_timeModule_CIN.py:728-744 vs :867-878), which is why the
    reference needs strategy='ddp_find_unused_parameters_true'.  (2) Convolution biases in front of an instance-norm CIN
    (cin.conv_before_film): the convolution runs without them, their reference gradient is round-off noise."""
    from . import cin as _cin
    n = 0
    for p in model.decoder_inv[model.N_layers].parameters():
        p.requires_grad_(False); n += p.numel()
    if _cin.SKIP_DEAD_BIAS:
        inst = lambda f: isinstance(getattr(f, 'norm', None), nn.InstanceNorm2d) and not f.norm.track_running_stats
        for m in model.modules():
            pairs = []
            if isinstance(m, cno_model.CNOBlock) and m.is_time:
                pairs = [(m.convolution, m.in_norm_conditiner)]
            elif isinstance(m, cno_model.ResidualBlock) and m.is_time:
                pairs = [(m.convolution1, m.in_norm_conditiner1), (m.convolution2, m.in_norm_conditiner2)]
            for conv, film in pairs:
                if conv.bias is not None and inst(film) and conv.bias.requires_grad:
                    conv.bias.requires_grad_(False); n += conv.bias.numel()
    return n


GATHER_ON_SIDE_STREAM = os.environ.get('CNO_GATHER_ON_SIDE_STREAM', '1') == '1'


class Trainer:
    """One optimisation step = forward, relative L1 (CNO_timeModule_CIN.py:939), backward, (all-reduce), AdamW (:996).

    Data parallelism: batch sharding, the gradient all-reduce is the only exchange.  With more than one rank the
    gradients travel through ONE flat fp32 buffer cut into `buckets` contiguous ranges (reverse registration order = the order
    backward produces them).  A post-accumulate hook counts the parameters of a bucket; when the last one has its gradient the
    bucket is gathered into the flat buffer with one multi-tensor copy, `p.grad` of its parameters become views of that range
    (AdamW then reads the reduced values in place: no copy back, no scaling pass) and the bucket's `ncclAllReduce(AVG)` is issued
    on a side stream, so it overlaps the rest of backward.  Gradients start every step as None: autograd hands over its own
    buffers instead of adding into stale ones (an `add_` per parameter - 1,268 launches, 2.4 ms of the 31 ms step - when
    `p.grad` pre-exists, profiles/r2_train_step_breakdown.txt).  A single rank does none of this.  All of it is capturable: with
    `graph=True` the whole step (forward, backward, bucket copies + all-reduces on the forked stream, AdamW) replays as one
    CUDA graph per rank.  `mode='finetune'`: CNO_FineTune.py workload (finetune.initialize_FT wrappers, three AdamW groups)."""

    def __init__(self, cfg=None, device='cuda', ddp=False, graph=False, autocast=None, buckets=4, mode='train', ft=None, side_stream=True, optimizer='multi'):
        self.device = torch.device(device)
        if mode == 'finetune':
            from . import finetune
            ft = ft or finetune.wide_finetune_config()
            self.model = finetune.initialize_FT(cno_model.CNO_time(**ft['model']), **ft['ft']).to(self.device)
        else:
            cfg = dict(cfg or cno_model.train_config_L())
            self.model = cno_model.CNO_time(**cfg).to(self.device)
        freeze_dead_parameters(self.model)
        self.world = dist.get_world_size() if (ddp and dist.is_initialized()) else 1
        self.params = [p for p in self.model.parameters() if p.requires_grad]
        if self.world > 1:
            with torch.no_grad():                      # same start on every rank
                for p in self.model.parameters():
                    dist.broadcast(p, 0)
        # ---- flat gradient buffer, bucketed (multi-rank only) -------------------------------------------------------
        self.flat_grad, self.buckets, self._bucket_of, self._views = None, [], {}, {}
        self.comm_stream, self.dp_mode = None, None
        if self.world > 1:
            order = self.params[::-1]
            total = sum(p.numel() for p in order)
            self.flat_grad = torch.zeros(total, dtype=torch.float32, device=self.device)
            nb = max(1, min(int(buckets), len(order)))
            target, o, start, members = total / nb, 0, 0, []
            for p in order:
                self._views[p] = self.flat_grad[o:o + p.numel()].as_strided(p.shape, p.stride())     # same layout as the parameter
                self._bucket_of[p] = len(self.buckets)
                members.append(p)
                o += p.numel()
                if o - start >= target and len(self.buckets) < nb - 1:
                    self.buckets.append((start, o, members)); start, members = o, []
            self.buckets.append((start, o, members))
            self._pending = [len(m) for _, _, m in self.buckets]
            self._launched = [False] * len(self.buckets)
            if self.device.type == 'cuda':
                self.comm_stream = torch.cuda.Stream(self.device)
            self.dp_mode = f'flat buffer, {len(self.buckets)} buckets, all-reduce(AVG) overlapped with backward'
            for p in order:
                p.register_post_accumulate_grad_hook(self._grad_ready)
        if mode == 'finetune':
            from . import finetune
            groups = finetune.finetune_param_groups(self.model, **ft['lrs'])
            wd = ft['model']['weight_decay']
        else:
            groups, wd = [{'params': self.params, 'lr': self.model.lr}], self.model.weight_decay
        if self.device.type == 'cuda' and optimizer == 'multi':
            from .optim import MultiTensorAdamW          # every parameter tensor of every group in ONE launch (csrc/adamw.cu)
            self.opt = MultiTensorAdamW(groups, lr=self.model.lr, weight_decay=wd)
        else:                                            # optimizer='torch': torch's fused multi-tensor AdamW (~32 launches, 0.8 ms)
            self.opt = torch.optim.AdamW(groups, lr=self.model.lr, weight_decay=wd, capturable=bool(graph and self.device.type == 'cuda'),
                                         **(dict(fused=True) if self.device.type == 'cuda' else {}))
        # FILM parameter gradients on a side stream (cin.PARAM_GRAD_STREAM): joined before they are gathered / used by AdamW
        self._side = None
        if self.device.type == 'cuda' and side_stream:
            from . import cin as _cin
            if _cin.PARAM_GRAD_STREAM is None or _cin.PARAM_GRAD_STREAM.device != self.device:
                _cin.PARAM_GRAD_STREAM = torch.cuda.Stream(self.device)
            self._side = _cin
        self.autocast = autocast
        self.graph = None
        self.use_graph = bool(graph) and self.device.type == 'cuda'
        self.static = None

    # ---- gradient exchange -------------------------------------------------------------------------------------------
    def _launch_bucket(self, b):
        start, end, members = self.buckets[b]
        self._launched[b] = True
        have = [p for p in members if p.grad is not None]
        side = self._side.PARAM_GRAD_STREAM if (self._side is not None and self.comm_stream is not None and GATHER_ON_SIDE_STREAM) else None
        if side is None and self._side is not None:
            self._side.join_param_grad_stream()
        cur = torch.cuda.current_stream(self.device) if self.comm_stream is not None else None
        if side is not None:
            # Most gradients of a bucket (convolution weights, FILM parameters) are produced on the parameter-gradient side stream.  The
            # gather runs THERE, behind them, so the compute stream never waits for weight gradients in the middle of backward; the few
            # gradients produced on the compute stream are covered by the wait below and handed over with record_stream.
            side.wait_stream(cur)
        with (torch.cuda.stream(side) if side is not None else contextlib.nullcontext()):
            if len(have) < len(members):               # a parameter without a gradient this step contributes zeros
                self.flat_grad[start:end].zero_()
            if have:
                grads = [p.grad for p in have]
                torch._foreach_copy_([self._views[p] for p in have], grads)
                if side is not None:
                    for g in grads:
                        g.record_stream(side)
        for p in members:
            p.grad = self._views[p]
        if self.comm_stream is None:                   # CPU / gloo: same arithmetic, no overlap
            dist.all_reduce(self.flat_grad[start:end])
            self.flat_grad[start:end].mul_(1.0 / self.world)
            return
        self.comm_stream.wait_stream(side if side is not None else cur)      # the bucket is complete on that stream
        with torch.cuda.stream(self.comm_stream):
            dist.all_reduce(self.flat_grad[start:end], op=dist.ReduceOp.AVG)

    def _grad_ready(self, p):
        b = self._bucket_of[p]
        self._pending[b] -= 1
        if self._pending[b] == 0 and not self._launched[b]:
            self._launch_bucket(b)

    def _finish_allreduce(self):
        for b in range(len(self.buckets)):             # buckets holding a parameter that got no gradient this step
            if not self._launched[b]:
                self._launch_bucket(b)
        if self.comm_stream is not None:
            torch.cuda.current_stream(self.device).wait_stream(self.comm_stream)
        self._pending = [len(m) for _, _, m in self.buckets]
        self._launched = [False] * len(self.buckets)

    def grad_vector(self):
        """all gradients of the last step as one vector (reverse registration order); test / diagnostic helper"""
        if self.flat_grad is not None:
            return self.flat_grad.clone()
        return torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in self.params[::-1]])

    # ---- step --------------------------------------------------------------------------------------------------------
    def _step_eager(self, t, x, y):
        for p in self.params:
            p.grad = None                               # autograd hands over its buffers: no add_ into stale gradients
        with torch.autocast(self.device.type, dtype=self.autocast, enabled=self.autocast is not None):
            pred = self.model(x, t)
        loss = self.model.training_loss(pred.float(), y.float())
        loss.backward()
        if self._side is not None:
            self._side.join_param_grad_stream()
        if self.world > 1:
            self._finish_allreduce()
        self.opt.step()
        return loss.detach()

    def step(self, t, x, y):
        if not self.use_graph:
            return self._step_eager(t, x, y)
        if self.graph is None:
            self.static = (t.clone(), x.clone(), y.clone())
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                for _ in range(3):                       # warm-up outside capture (filter taps cached, cuDNN plans chosen)
                    self._step_eager(*self.static)
            torch.cuda.current_stream().wait_stream(s)
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self.static_loss = self._step_eager(*self.static)
        for dst, src in zip(self.static, (t, x, y)):
            dst.copy_(src, non_blocking=True)
        self.graph.replay()
        return self.static_loss

    def close(self):
        """Drop the captured graph BEFORE the process group goes away: a CUDA graph that holds NCCL kernels keeps the
        communicator's resources referenced, and destroying the communicator first blocks in its destructor."""
        if self.device.type == 'cuda':
            torch.cuda.synchronize(self.device)
        self.graph = None
        self.static = None
        self.static_loss = None
        if self._side is not None:
            self._side.PARAM_GRAD_STREAM = None
            self._side._side_keepalive.clear()
            self._side = None
        if self.device.type == 'cuda':
            torch.cuda.synchronize(self.device)


def synthetic_finetune_batch(batch, ft, device='cuda', seed=0, dtype=torch.float32):
    """fine-tuning sample shapes: x[B, new_in, S, S], t[B], y[B, new_out, S_out, S_out].  Band-limited random fields (coarse
    noise, bicubically up-sampled: the alias-free activations pass them) and a target that depends on the input (the first
    output channels of x at the output size, mixed with the lead time), so the loss has something to learn."""
    g = torch.Generator(device='cpu').manual_seed(seed)
    S, So = ft['in_size'], ft['out_size']
    coarse = torch.randn(batch, ft['in_dim'], max(S // 8, 2), max(S // 8, 2), generator=g)
    x = torch.nn.functional.interpolate(coarse, size=(S, S), mode='bicubic', align_corners=False)
    t = torch.randint(1, 8, (batch,), generator=g).float() / 7.0
    src = x[:, [i % ft['in_dim'] for i in range(ft['out_dim'])]]
    y = torch.nn.functional.interpolate(src, size=(So, So), mode='bilinear', align_corners=False, antialias=True)
    y = y * (0.5 + t.view(-1, 1, 1, 1))
    return t.to(device), x.contiguous().to(device, dtype), y.contiguous().to(device, dtype)
