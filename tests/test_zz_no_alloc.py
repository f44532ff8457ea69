"""cno_cin_fwd_noalloc (include/cno_b200.h): the CIN forward for callers that cannot allocate device memory.  It has no counterpart
in the reference; its per-plane statistics live in a fixed array inside the library's device image instead of caller buffers.
Checked through the C ABI against cno_cin_fwd on the same inputs, on calls in a row that share (and overwrite) that array.
Named test_zz_* so that pytest collects it last: a failure here must not hide other tests."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = 'cuda'
MAX_PLANES = 65536                   # CIN_NOALLOC_MAX_PLANES in csrc/cin.cu, stated in include/cno_b200.h
DT = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}


@pytest.fixture(scope='module')
def cno():
    import pregenerating_pde_b200 as pkg
    assert torch.cuda.is_available()
    pkg._lib.load()
    return pkg


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def both(cno, N, C, H, W, dtype, seed):
    """(y of cno_cin_fwd with caller statistics buffers, y of cno_cin_fwd_noalloc) on the same inputs"""
    from pregenerating_pde_b200.torch_utils.custom_ops import _ptr, _stream
    lib = cno._lib.load()
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(N, C, H, W, generator=g) * 1.7 + 0.3).to(dtype).to(DEV)
    gamma, beta = torch.randn(C, generator=g).to(DEV), torch.randn(C, generator=g).to(DEV)
    scale, shift = torch.randn(N, C, generator=g).to(DEV), torch.randn(N, C, generator=g).to(DEV)
    y_ref, y_new = torch.empty_like(x), torch.full_like(x, float('nan'))
    mean, rstd = torch.empty(N, C, device=DEV), torch.empty(N, C, device=DEV)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = lib.cno_cin_fwd(_ptr(x), _ptr(y_ref), _ptr(gamma), _ptr(beta), _ptr(scale), _ptr(shift), _ptr(mean), _ptr(rstd),
                         DT[dtype], N, C, H * W, 1e-5, _stream(x))
    assert rc == 0, lib.cno_last_error()
    rc = lib.cno_cin_fwd_noalloc(_ptr(x), _ptr(y_new), _ptr(gamma), _ptr(beta), _ptr(scale), _ptr(shift),
                                 DT[dtype], N, C, H * W, 1e-5, _stream(x))
    assert rc == 0, lib.cno_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return y_ref, y_new


# Every launch path of the forward: planes in registers (8^2 .. 128^2), the thread-block cluster (256^2), the streaming kernel,
# which writes the statistics and reads them back (ragged 10x10 / 37x53 planes: no 16-byte rows), and 16-bit storage.
# Consecutive cases differ in N*C, so a call sees the previous call's statistics in the shared array.
CALLS = [(32, 256, 8, 8, torch.float32), (2, 3, 10, 10, torch.float32), (4, 16, 128, 128, torch.float32), (3, 5, 37, 53, torch.float32),
         (1, 2, 256, 256, torch.float32), (2, 7, 37, 53, torch.bfloat16), (4, 8, 64, 64, torch.bfloat16), (2, 4, 32, 32, torch.float16)]


def test_cin_fwd_noalloc_matches_cin_fwd_on_calls_in_a_row(cno):
    for rounds in range(2):                      # the whole sequence twice: every case also runs on its own stale statistics
        for i, (N, C, H, W, dtype) in enumerate(CALLS):
            y_ref, y_new = both(cno, N, C, H, W, dtype, seed=100 * rounds + i)
            assert not torch.isnan(y_new.float()).any()
            tol = 1e-5 if dtype == torch.float32 else 2e-2          # the tolerances of tests/test_gpu_parity.py
            assert rel(y_new, y_ref) < tol, (rounds, N, C, H, W, dtype, rel(y_new, y_ref))
            assert torch.equal(y_new, y_ref)     # same kernels, same arithmetic: in fact bit-identical


def test_cin_fwd_noalloc_same_input_twice(cno):
    """two calls in a row on the same input give the same output (nothing carried over between calls matters)"""
    a_ref, a = both(cno, 3, 5, 37, 53, torch.float32, seed=7)
    b_ref, b = both(cno, 3, 5, 37, 53, torch.float32, seed=7)
    assert rel(a, a_ref) < 1e-5 and rel(b, b_ref) < 1e-5
    assert torch.equal(a, b)


def test_cin_fwd_noalloc_rejects_sizes_beyond_the_scratch(cno):
    from pregenerating_pde_b200.torch_utils.custom_ops import _ptr, _stream
    lib = cno._lib.load()
    d = torch.zeros(16, device=DEV)              # never touched: the size check comes before any launch
    for n, c in [(1, MAX_PLANES + 1), (257, 256), (MAX_PLANES, 2)]:
        rc = lib.cno_cin_fwd_noalloc(_ptr(d), _ptr(d), _ptr(d), _ptr(d), _ptr(d), _ptr(d), 0, n, c, 4, 1e-5, _stream(d))
        assert rc != 0
        msg = lib.cno_last_error().decode()
        assert 'exceed' in msg and str(MAX_PLANES) in msg, msg
    assert lib.cno_cin_fwd_noalloc(None, None, None, None, None, None, 0, 1, 1, 4, 1e-5, None) != 0
    # the largest accepted size runs (n * c == the limit; 2x2 planes)
    y_ref, y_new = both(cno, 256, 256, 2, 2, torch.float32, seed=3)
    assert torch.equal(y_new, y_ref)
    torch.cuda.synchronize()
