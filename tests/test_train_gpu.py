"""GPU tests of the training harness: the CUDA-graph-captured step equals the eager step, the fine-tune
workload (BASELINE.json configs[4] shape, shrunk) runs under bf16 autocast, and the captured forward-only rollout loop of
TestCNO_ALL.py:124-188 equals the eager loop."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = 'cuda'
SMALL = dict(in_dim=7, in_size=32, N_layers=2, N_res=1, N_res_neck=2, channel_multiplier=8, out_dim=3, is_time=1, nl_dim=[2, 3],
             latent_lift_proj_dim=16, lr=1e-3, weight_decay=1e-6)


@pytest.fixture(scope='module')
def cno():
    import pregenerating_pde_b200 as pkg
    pkg._lib.load()
    return pkg


def _params(tr):
    return torch.cat([p.detach().reshape(-1) for p in tr.model.parameters()])


def test_graph_step_equals_eager_step(cno):
    """Same parameters, same batches: the captured step must produce the gradients and losses of the eager step.
    Parameters after several AdamW steps are NOT compared element-wise: convolution biases in front of an instance norm have
    a mathematically zero gradient, so their computed gradient is fp32 round-off (and the order of the db atomics), which
    AdamW's m / sqrt(v) turns into +-lr steps of arbitrary sign.  Gradients are compared with the optimiser frozen (lr = 0),
    the training trajectory through its losses."""
    torch.manual_seed(0)
    cfg0 = dict(SMALL, lr=0.0, weight_decay=0.0)
    batches = [cno.train.synthetic_fpo_batch(4, size=32, device=DEV, seed=s) for s in range(3)]
    # TF32 off for the gradient comparison: with TF32 convolutions cuDNN's algorithm choice (which may differ between the eager and the
    # captured launches) moves gradients by ~1e-4 of their maximum, i.e. the comparison would sit on its own tolerance
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        a = cno.train.Trainer(cfg0, device=DEV, graph=False)
        b = cno.train.Trainer(cfg0, device=DEV, graph=True)
        b.model.load_state_dict(a.model.state_dict())
        for bt in batches + batches:
            la, ga = float(a.step(*bt)), a.grad_vector()
            lb, gb = float(b.step(*bt)), b.grad_vector()
            assert torch.isfinite(gb).all()
            assert abs(la - lb) <= 1e-5 * abs(la), (la, lb)
            # fp32 floor: the db atomics and cuDNN's algorithm choice may differ between the warm-up and the captured launches
            assert float((ga - gb).abs().max() / ga.abs().max()) < 1e-4
        b.close()
    finally:
        torch.backends.cudnn.allow_tf32 = tf32
    # with the optimiser running: same loss trajectory (3 warm-up steps precede the capture, give the eager trainer the same history)
    torch.manual_seed(0)
    a = cno.train.Trainer(SMALL, device=DEV, graph=False)
    b = cno.train.Trainer(SMALL, device=DEV, graph=True)
    b.model.load_state_dict(a.model.state_dict())
    for _ in range(3):
        a.step(*batches[0])
    la = [float(a.step(*bt)) for bt in batches + batches]
    lb = [float(b.step(*bt)) for bt in batches + batches]
    assert all(abs(x - y) / abs(x) < 1e-3 for x, y in zip(la, lb)), (la, lb)
    assert torch.isfinite(_params(b)).all()
    b.close()


def test_finetune_step_bf16_autocast(cno):
    ft = dict(model=dict(in_dim=5, in_size=64, N_layers=2, N_res=1, N_res_neck=1, channel_multiplier=8, out_dim=4, is_time=1, nl_dim=[2, 3],
                         latent_lift_proj_dim=16, lr=1e-3, weight_decay=1e-6),
              ft=dict(old_in_dim=5, new_in_dim=4, new_out_dim=4, old_out_dim=4), lrs=dict(lr=1e-3, lr_emb=1e-2, lr_norm=1e-2),
              in_size=64, out_size=64, in_dim=4, out_dim=4)
    tr = cno.train.Trainer(device=DEV, mode='finetune', ft=ft, autocast=torch.bfloat16, graph=True)
    t, x, y = cno.train.synthetic_finetune_batch(4, ft, device=DEV)
    losses = [float(tr.step(t, x, y)) for _ in range(30)]
    assert all(l == l for l in losses) and min(losses[-5:]) < 0.9 * losses[0], losses      # trains: the loss on a fixed batch goes down
    assert len(tr.opt.param_groups) == 3
    tr.close()


def test_rollout_graph_equals_eager(cno):
    """forward-only autoregressive rollout (TestCNO_ALL.py:124-188: pred = model(inp, t); inp[:, :dim] = pred) captured as one
    CUDA graph of K model calls; no sign masks are written under no_grad."""
    torch.manual_seed(1)
    model = cno.cno_model.CNO_time(**SMALL).to(DEV).eval()
    t, x, _ = cno.train.synthetic_fpo_batch(2, size=32, device=DEV, seed=4)

    def rollout(inp, K=4):
        outs = []
        inp = inp.clone()
        for _ in range(K):
            pred = model(inp, t)
            inp = torch.cat([pred, inp[:, 3:]], 1)
            outs.append(pred)
        return torch.stack(outs)
    with torch.no_grad():
        ref = rollout(x)
        s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            rollout(x)
        torch.cuda.current_stream().wait_stream(s)
        g = torch.cuda.CUDAGraph()
        static_x = x.clone()
        with torch.cuda.graph(g):
            out = rollout(static_x)
        static_x.copy_(x)
        g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, ref)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs two GPUs')
def test_multi_rank_gradients_side_stream_gather(cno):
    """torchrun, 2 ranks over NCCL: reduced gradients with the bucket gather on the parameter-gradient side stream equal those with the
    gather on the compute stream, eager and captured, and are identical on every rank (tools/check_ddp_gather.py)"""
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr', '127.0.0.1',
                          '--master-port', '29541', os.path.join(root, 'tools', 'check_ddp_gather.py')], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and 'OK, worst relative gradient difference' in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]
