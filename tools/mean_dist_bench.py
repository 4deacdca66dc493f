import sys, torch, numpy as np
sys.path.insert(0,'/root/repo')
import uuo_mocap_b200
from uuo_mocap_b200 import ops
F,M,V=5000,45,6890
v=torch.randn(F,V,3,device='cuda'); m=torch.randn(F,M,3,device='cuda')
for _ in range(2): ops.mean_distance_argmin(m,v)
torch.cuda.synchronize()
e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): ops.mean_distance_argmin(m,v,finalize=False)
e1.record(); torch.cuda.synchronize(); print("mean_dist ms", e0.elapsed_time(e1)/5)
