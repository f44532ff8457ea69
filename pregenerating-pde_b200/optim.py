"""AdamW for the CNO training loop (TrainCNO_time_L.py:185-196, CNO_timeModule_CIN.py:996, CNO_FineTune.py three-group set-up) as ONE
kernel launch over all parameter tensors of all groups (csrc/adamw.cu, `cno_adamw_multi`).  Same arithmetic and the same constructor
arguments as `torch.optim.AdamW` (amsgrad / maximize are not supported and not used by the reference).

The kernel reads a table of (param, grad, exp_avg, exp_avg_sq, numel, lr, weight_decay) records from device memory.  Gradients of a step
are whatever `p.grad` holds when `step()` is called: the table is rebuilt (one small pinned-memory upload) whenever a pointer, a learning
rate or a weight decay changed since the last call.  Under CUDA-graph capture the upload is captured with the step, and the captured
gradient buffers keep their addresses, so replays need no host work at all.  The step count lives on the device."""
import numpy as np
import torch

from . import _lib
from .torch_utils.custom_ops import _ptr, _stream

_RECORD = np.dtype([('param', '<u8'), ('grad', '<u8'), ('exp_avg', '<u8'), ('exp_avg_sq', '<u8'), ('numel', '<i8'),
                    ('lr', '<f4'), ('weight_decay', '<f4'), ('first_chunk', '<i4'), ('pad', '<i4')])     # CnoAdamTensor, include/cno_b200.h
assert _RECORD.itemsize == 56


def _dense(t):
    """every element of the storage range exactly once (contiguous in some dimension order)"""
    if t.is_contiguous():
        return True
    expect = 1
    for n, st in sorted(((n, st) for n, st in zip(t.shape, t.stride()) if n > 1), key=lambda v: v[1]):
        if st != expect:
            return False
        expect *= n
    return True


def _same_layout(a, b):
    """same strides on every dimension that has more than one element"""
    return a.shape == b.shape and all(sa == sb for n, sa, sb in zip(a.shape, a.stride(), b.stride()) if n > 1)


class MultiTensorAdamW(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2):
        if not (0.0 <= betas[0] < 1.0 and 0.0 <= betas[1] < 1.0):
            raise ValueError(f'invalid betas {betas}')
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self._params = [p for g in self.param_groups for p in g['params']]
        if not self._params:
            raise ValueError('no parameters')
        dev = self._params[0].device
        for g in self.param_groups:
            if tuple(g['betas']) != tuple(betas) or g['eps'] != eps:
                raise ValueError('one (betas, eps) pair for all groups')          # one launch, one set of scalars
            for p in g['params']:
                if p.dtype != torch.float32 or p.device != dev or not _dense(p) or dev.type != 'cuda':
                    raise ValueError('MultiTensorAdamW: dense fp32 CUDA parameters on one device')
        self._lib = _lib.load()                                                  # raises without the CUDA library: no fallback
        self._chunk = int(self._lib.cno_adamw_chunk())
        total = sum(p.numel() for p in self._params)
        # moments: one allocation each, a view per parameter (state_dict-compatible with torch.optim.AdamW: exp_avg, exp_avg_sq, step)
        self._m, self._v = torch.zeros(total, device=dev), torch.zeros(total, device=dev)
        self._step = torch.zeros(1, device=dev)
        o, first, chunk_tensor = 0, 0, []
        self._first_chunk = []
        for i, p in enumerate(self._params):
            n = p.numel()
            # the update is elementwise over storage: moments (and gradients) must share the parameter's strides (e.g. channels_last weights)
            self.state[p] = dict(step=self._step, exp_avg=self._m[o:o + n].as_strided(p.shape, p.stride()),
                                 exp_avg_sq=self._v[o:o + n].as_strided(p.shape, p.stride()))
            nch = (n + self._chunk - 1) // self._chunk
            self._first_chunk.append(first)
            chunk_tensor += [i] * nch
            first += nch
            o += n
        self._n_chunks = first
        self._chunk_tensor = torch.tensor(chunk_tensor, dtype=torch.int32, device=dev)
        self._table_dev = torch.empty(len(self._params) * _RECORD.itemsize, dtype=torch.uint8, device=dev)
        # pinned staging for the table, allocated up front (no allocation may happen under stream capture): a ring for eager steps and
        # dedicated buffers for captured steps, which a graph re-reads on every replay and which are therefore never written again
        nbytes = len(self._params) * _RECORD.itemsize
        self._ring = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(4)]
        self._capture_bufs = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(2)]
        self._ring_pos, self._key, self._captured = 0, None, []

    def _table(self):
        """device table for the current gradients; uploaded only when something changed"""
        rec = np.zeros(len(self._params), dtype=_RECORD)
        i = 0
        for g in self.param_groups:
            for p in g['params']:
                gr = p.grad
                if gr is not None and (gr.dtype != torch.float32 or gr.is_sparse or not _same_layout(gr, p)):
                    raise RuntimeError('MultiTensorAdamW: dense fp32 gradients with the strides of their parameter')
                st = self.state[p]
                rec[i] = (p.data_ptr(), 0 if gr is None else gr.data_ptr(), st['exp_avg'].data_ptr(), st['exp_avg_sq'].data_ptr(), p.numel(),
                          float(g['lr']), float(g['weight_decay']), self._first_chunk[i], 0)
                i += 1
        key = rec.tobytes()
        capturing = torch.cuda.is_current_stream_capturing()
        if key != self._key or capturing:
            if capturing:
                if not self._capture_bufs:
                    raise RuntimeError('MultiTensorAdamW: more than two captured steps per optimiser')
                host = self._capture_bufs.pop()
                self._captured.append(host)                                       # the graph's copy node reads it on every replay: keep it alive
            else:
                host = self._ring[self._ring_pos]
                self._ring_pos = (self._ring_pos + 1) % len(self._ring)
                if self._ring_pos == 0:
                    torch.cuda.current_stream(self._step.device).synchronize()    # the oldest upload has certainly left its buffer
            host.numpy()[:] = rec.view(np.uint8)
            self._table_dev.copy_(host, non_blocking=True)
            self._key = None if capturing else key                               # after a capture the next eager call uploads again
        return self._table_dev

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        g0 = self.param_groups[0]
        table = self._table()
        rc = self._lib.cno_adamw_multi(_ptr(table), _ptr(self._chunk_tensor), self._n_chunks, _ptr(self._step),
                                       float(g0['betas'][0]), float(g0['betas'][1]), float(g0['eps']), _stream(self._step))
        _lib.check(rc, 'cno_adamw_multi')
        return loss
