"""CNO training step harness: the workload of TrainCNO_time_L.py (CNO_Experiments/TrainCNO_time_L.py:167-224 builds
the model and hands it to a Lightning Trainer with DDP) restated without Lightning: synthetic FPO-shaped batches,
forward + relative-L1 loss + backward + AdamW, one process per GPU with torch DDP over NCCL (gradient all-reduce is
the only exchange), optional whole-step CUDA-graph capture on a single GPU.
"""
import math
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


def freeze_dead_parameters(model):
    """decoder_inv[N_layers] is constructed but never called (CNO_timeModule_CIN.py:728-744 vs :867-878), which is why
    the reference needs strategy='ddp_find_unused_parameters_true'.  Parameters without a gradient are skipped by
    AdamW either way, so marking them frozen changes nothing numerically and lets DDP run without the search."""
    dead = model.decoder_inv[model.N_layers]
    n = 0
    for p in dead.parameters():
        p.requires_grad_(False); n += p.numel()
    return n


class Trainer:
    """One optimisation step = forward, relative L1 (CNO_timeModule_CIN.py:939), backward, (all-reduce), AdamW (:996)."""

    def __init__(self, cfg=None, device='cuda', ddp=False, graph=False, autocast=None, channels_last=False):
        self.device = torch.device(device)
        cfg = dict(cfg or cno_model.train_config_L())
        self.model = cno_model.CNO_time(**cfg).to(self.device)
        freeze_dead_parameters(self.model)
        self.net = self.model
        self.world = dist.get_world_size() if (ddp and dist.is_initialized()) else 1
        # data parallelism, two flavours with identical numerics (mean of the per-rank gradients):
        #  'ddp'  : torch DistributedDataParallel (bucketed all-reduce overlapped with backward), eager launches
        #  'flat' : one NCCL all-reduce of the flattened gradients after backward; capturable, so the whole step
        #           (forward, backward, all-reduce, AdamW) replays as ONE CUDA graph per rank
        self.dp_mode = None
        if ddp:
            self.dp_mode = 'flat' if (graph and self.device.type == 'cuda') else 'ddp'
        if self.dp_mode == 'ddp':
            self.net = nn.parallel.DistributedDataParallel(self.model, device_ids=[self.device.index] if self.device.type == 'cuda' else None,
                                                           broadcast_buffers=False,
                                                           gradient_as_bucket_view=True, bucket_cap_mb=100, static_graph=True)
        elif self.dp_mode == 'flat':
            with torch.no_grad():                      # same start on every rank (DDP does this broadcast itself)
                for p in self.model.parameters():
                    dist.broadcast(p, 0)
        self.opt = torch.optim.AdamW([p for p in self.model.parameters() if p.requires_grad], lr=self.model.lr,
                                     weight_decay=self.model.weight_decay, capturable=bool(graph and self.dp_mode != 'ddp'),
                                     **(dict(fused=True) if self.device.type == 'cuda' else {}))    # one multi-tensor kernel family
        self.autocast = autocast
        self.graph = None
        self.use_graph = bool(graph) and self.dp_mode != 'ddp' and self.device.type == 'cuda'
        self.static = None

    def _step_eager(self, t, x, y):
        self.opt.zero_grad(set_to_none=True)
        with torch.autocast(self.device.type, dtype=self.autocast, enabled=self.autocast is not None):
            pred = self.net(x, t)
        loss = self.model.training_loss(pred.float(), y.float())
        loss.backward()
        if self.dp_mode == 'flat':
            self._allreduce_flat()
        self.opt.step()
        return loss.detach()

    def _allreduce_flat(self):
        grads = [p.grad for p in self.model.parameters() if p.grad is not None]
        flat = torch.cat([g.reshape(-1) for g in grads])
        dist.all_reduce(flat)
        flat.mul_(1.0 / self.world)
        torch._foreach_copy_(grads, [c.view_as(g) for c, g in zip(flat.split([g.numel() for g in grads]), grads)])

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
            self.opt.zero_grad(set_to_none=True)
            with torch.cuda.graph(self.graph):
                self.static_loss = self._step_eager(*self.static)
        for dst, src in zip(self.static, (t, x, y)):
            dst.copy_(src, non_blocking=True)
        self.graph.replay()
        return self.static_loss
