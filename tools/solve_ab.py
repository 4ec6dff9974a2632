import sys, json, torch
sys.path.insert(0,'/root/repo')
from opticalflow_b200 import engine, _lib, synthetic
import numpy as np
dev=torch.device('cuda',0)
B=16
movie=synthetic.rotating_shearing_movie(1024,1024,B+1,seed=2).astype(np.float32)
fr=engine.blur(torch.from_numpy(movie).to(dev),2.5)
P,C=fr[:-1].contiguous(),fr[1:].contiguous()
b=engine.build_rhs(P,C)
a=torch.full((B,),1e5,dtype=torch.float64,device=dev); be=torch.full((B,),1e4,dtype=torch.float64,device=dev)
MODES = {'default': 0, 'no-half-step-exit': _lib.FLAG_NO_HALF_STEP_EXIT, 'pdl': _lib.FLAG_PDL, 'strips': _lib.FLAG_STRIP_ITEMS,
         'serial-setup': _lib.FLAG_SERIAL_SETUP}
names = sys.argv[1].split(',') if len(sys.argv) > 1 else ['default', 'serial-setup', 'default', 'serial-setup']
for name,flags in [(n, MODES[n]) for n in names]:
    opt=engine.make_options('mg',flags=flags)
    ts=[]
    for r in range(6):
        x=torch.zeros_like(b)
        torch.cuda.synchronize()
        e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        e0.record(); it,rr,cv=engine.solve(P,b,x,a,be,1e-6,1000,opt); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    print(json.dumps(dict(mode=name, solve_ms=min(ts[1:]), iters=float(it.float().mean()), conv=int(cv.sum()), xnorm=float(x.double().norm()))))
