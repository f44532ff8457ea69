"""graph-timed fused filtered_lrelu fwd/bwd for one shape (uses bench_sweep helpers): N C S_in S_out [dtype]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench_sweep as B
a = sys.argv[1:]
rows = []
dt = {'f32': torch.float32, 'bf16': torch.bfloat16}[a[4] if len(a) > 4 else 'f32']
B.sweep_filtered_lrelu(rows, int(a[0]), int(a[1]), int(a[2]), int(a[3]), dt, 10)
print(' | '.join(f"{r['shape']} {r['dir']} {r['us']:.1f} us {r['gbs']:.0f} GB/s" for r in rows))
