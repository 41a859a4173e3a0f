import sys; sys.path.insert(0,'/root/repo')
import torch, numpy as np
from ekat_b200.lin_interp import LinInterp, fill
dev=torch.device('cuda:0')
for (ncol,km1,km2,N) in [(21600,72,128,1),(5,72,128,1),(10,50,50,1),(10,37,36,16)]:
    x1,y1,x2=fill(torch.float64,ncol,km1,km2,N,device=dev)
    li=LinInterp(ncol,km1,km2,torch.float64,N,device=dev)
    y2=torch.zeros((ncol,li.pitch2),dtype=torch.float64,device=dev)
    for nm,f in (('run',lambda: li.run(x1,x2,y1,y2)),('setup',lambda: li.setup(x1,x2)),('apply',lambda: li.lin_interp(x1,x2,y1,y2))):
        try:
            f(); torch.cuda.synchronize(); print(nm,'ok',ncol,km1,km2,N)
        except Exception as e: print(nm,'FAIL',ncol,km1,km2,N,e)
