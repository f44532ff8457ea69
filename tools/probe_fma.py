"""FP32 issue-rate probe on the B200: scalar FFMA vs packed FFMA2 (fma.rn.f32x2), per SM.
Writes gpurun_out/probe_fma.json.  Design input for the fused FIR kernel."""
import ctypes, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pregenerating_pde_b200 as cno

lib = cno._lib.load()
out = torch.empty(148 * 8 * 256, device='cuda')
res = {}
for variant, name in ((0, 'ffma'), (1, 'ffma2')):
    for blocks_per_sm in (2, 4, 8):
        blocks, iters = 148 * blocks_per_sm, 4000
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        lib.cno_probe_fma(ctypes.c_void_p(out.data_ptr()), variant, blocks, 256, 10, st)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); lib.cno_probe_fma(ctypes.c_void_p(out.data_ptr()), variant, blocks, 256, iters, st); e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        fma = blocks * 256 * iters * 8 * 8 * 2      # scalar FMAs (each FFMA2 counts as two)
        res[f'{name}_{blocks_per_sm}cta'] = dict(ms=round(ms, 3), tflops=round(2 * fma / ms / 1e9, 2),
                                                 fma_per_clk_per_sm_at_1965MHz=round(fma / (ms * 1e-3) / 148 / 1.965e9, 1))
os.makedirs('gpurun_out', exist_ok=True)
json.dump(res, open('gpurun_out/probe_fma.json', 'w'), indent=1)
print(json.dumps(res, indent=1))
