"""Ten-second check of the tcgen05 gradient call at 128 / 1,000 / 65,536 rows (finite sums, wall time): run it first under
`timeout 45` after touching csrc/mlp_tc.cu -- a kernel that hangs costs GPU minutes, this costs seconds."""
import sys, time
sys.path.insert(0, "deep-learning-at-scale_b200"); sys.path.insert(0, "oracle")
import torch, dlas
from dlas import ops
DIMS=[32,256,256,1]
for n in (128, 1000, 65536):
    X,y=ops.synth_rows(n,32)
    p=torch.randn(ops.mlp_n_params(DIMS),device="cuda")*0.05
    ws=ops.mlp_workspace(DIMS,dlas.MODE_TC); gs=torch.zeros(p.numel()+1,device="cuda")
    t=time.time(); ops.mlp_grad(DIMS,0,dlas.MODE_TC,p,X,y,gs,ws); torch.cuda.synchronize()
    print(n, "finite", bool(torch.isfinite(gs).all()), "t=%.3f"%(time.time()-t), float(gs[-1]), flush=True)
