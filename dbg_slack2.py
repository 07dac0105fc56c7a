import sys, time
sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np
from __graft_entry__ import load, load_oracle
M,O=load(),load_oracle()
from parity import dense_scenarios
N,S=16,48
base=dict(k_hor=20,num_pieces=4,d_min=0.8,num_agents=N,slack_mode=1,slack_cost=50000.0,slack_decay=0.1)
cfg=M.default_config(max_scenarios=S,**base)
h=M.Handle(cfg)
st,tg=dense_scenarios(S,N,N,0.8,spread=1.15,vmax=1.0)
got=h.step_batch_host(st,tg)
its=h.last_iteration_counts(S*N*2).reshape(S,N,2)
for it in (0,1):
    x=its[:,:,it]
    print("it",it,"pct 10/50/90/99/100",np.percentile(x,[10,50,90,99,100]),"mean",x.mean(), "n>=20000",(x>=20000).sum())
    print("  mean by agent", x.mean(0).astype(int))
print(np.bincount(got['status'].ravel(),minlength=8))
