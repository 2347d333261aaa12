import sys, time, numpy as np, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from primate_b200.operators import MatrixFunction
from primate_b200.trace import hutch
from primate_b200.random import ProbeStream
from primate_b200 import _ffi
rng=np.random.default_rng(0); B=rng.standard_normal((1000,1000)); A=B@B.T/1000+0.1*np.eye(1000)
M=MatrixFunction(A,"log",deg=20,orth=0)
hutch(M,converge="count",count=200,seed=1)
for rep in range(3):
    torch.cuda.synchronize(); t0=time.perf_counter(); est=hutch(M,converge="count",count=200,seed=1); torch.cuda.synchronize(); print('hutch wall ms',1e3*(time.perf_counter()-t0))
S=ProbeStream(1000,"rademacher",seed=1)
for kb in (64,32):
    torch.cuda.synchronize(); l0=_ffi.lib().slq_launch_count(); e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    t0=time.perf_counter(); e0.record(); y=M.quad_range(S,0,200,kb); e1.record(); torch.cuda.synchronize()
    print('quad_range kb',kb,'wall ms',1e3*(time.perf_counter()-t0),'device ms',e0.elapsed_time(e1),'launches',_ffi.lib().slq_launch_count()-l0)
# one block timing pieces
X=S.block(64,0); op=M.device_op
torch.cuda.synchronize(); e0.record(); out=op.lanczos(X,20,1e-8,0,2); e1.record(); torch.cuda.synchronize(); print('lanczos block 64 device ms', e0.elapsed_time(e1))
Y=torch.empty_like(X)
e0.record()
for _ in range(20): op.matmat(X)
e1.record(); torch.cuda.synchronize(); print('dense matmat n=1000 k=64 ms each', e0.elapsed_time(e1)/20)
