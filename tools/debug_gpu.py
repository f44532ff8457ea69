"""GPU bring-up probe: runs each kernel family in its own subprocess so one fault does not hide the rest."""
import subprocess, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = {
 'fma0': "import ctypes;o=torch.zeros(148*256,device='cuda');print(lib.cno_probe_fma(ctypes.c_void_p(o.data_ptr()),0,148,256,10,None));torch.cuda.synchronize()",
 'fma1': "import ctypes;o=torch.zeros(148*256,device='cuda');print(lib.cno_probe_fma(ctypes.c_void_p(o.data_ptr()),1,148,256,10,None));torch.cuda.synchronize()",
 'bias_act': "x=torch.randn(4,8,16,16,device='cuda');b=torch.randn(8,device='cuda');y=cno.bias_act.bias_act(x,b,act='lrelu');torch.cuda.synchronize();print((y.cpu()-O.bias_act(x.cpu(),b.cpu(),act='lrelu')).abs().max())",
 'upfirdn': "x=torch.randn(2,3,16,16,device='cuda');f=torch.randn(12,device='cuda');y=cno.upfirdn2d.upfirdn2d(x,f,up=2,padding=[6,5,6,5]);torch.cuda.synchronize();print((y.cpu()-O.upfirdn2d(x.cpu(),f.cpu(),up=2,padding=[6,5,6,5])).abs().max())",
 'cin': "x=torch.randn(2,8,32,32,device='cuda');g=torch.randn(8,device='cuda');s=torch.randn(2,8,device='cuda');y=cno.cin.cin(x,g,g,s,s);torch.cuda.synchronize();print((y.cpu()-O.cin(x.cpu(),g.cpu(),g.cpu(),s.cpu(),s.cpu())).abs().max())",
 'flr_none': "cfg=O.lrelu_config(16,16);x=torch.randn(2,4,16,16,device='cuda');b=torch.randn(4,device='cuda');y=cno.filtered_lrelu.filtered_lrelu(x,cfg['fu'].cuda(),cfg['fd'].cuda(),b,up=2,down=2,padding=cfg['padding']);torch.cuda.synchronize();print((y.cpu()-O.filtered_lrelu(x.cpu(),cfg['fu'],cfg['fd'],b.cpu(),up=2,down=2,padding=cfg['padding'])).abs().max())",
 'flr_write': "cfg=O.lrelu_config(16,16);x=torch.randn(2,4,16,16,device='cuda').requires_grad_(True);b=torch.randn(4,device='cuda');y=cno.filtered_lrelu.filtered_lrelu(x,cfg['fu'].cuda(),cfg['fd'].cuda(),b,up=2,down=2,padding=cfg['padding']);torch.cuda.synchronize();print((y.detach().cpu()-O.filtered_lrelu(x.detach().cpu(),cfg['fu'],cfg['fd'],b.cpu(),up=2,down=2,padding=cfg['padding'])).abs().max())",
}
PRE = "import sys;sys.path.insert(0,%r);import torch;import pregenerating_pde_b200 as cno;from oracle import cno_oracle as O;lib=cno._lib.load();" % ROOT
if __name__ == '__main__':
    names = sys.argv[1:] or list(CASES)
    for n in names:
        r = subprocess.run([sys.executable, '-c', PRE + CASES[n]], capture_output=True, text=True, timeout=300)
        print(f'== {n}: rc={r.returncode} out={r.stdout.strip()[-200:]} err={r.stderr.strip()[-300:]}', flush=True)
