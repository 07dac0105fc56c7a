import sys
sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np, torch
from __graft_entry__ import load
M=load()
import bench
N,S=16,4096
cfg=M.default_config(k_hor=20,num_pieces=4,d_min=0.8,num_agents=N,max_scenarios=S,slack_mode=1,slack_cost=50000.0,slack_decay=0.1)
h=M.Handle(cfg)
st,tg=bench.make_swarm(S,N,7)
d_st,d_tg=torch.tensor(st,device="cuda"),torch.tensor(tg,device="cuda")
cp=torch.zeros((S,N,h.n),dtype=torch.float64,device="cuda"); stat=torch.zeros((S,N,2),dtype=torch.int32,device="cuda")
for _ in range(3): h.step_batch(S,d_st,d_tg,cp,stat)
torch.cuda.synchronize()
print("fallbacks",h.slack_fallback_count())
