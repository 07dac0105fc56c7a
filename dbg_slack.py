import sys, time
sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np
from __graft_entry__ import load, load_oracle
M,O=load(),load_oracle()
from parity import dense_scenarios
for N,S in ((2,6),(5,6),(8,4),(16,3),(16,256)):
    base=dict(k_hor=20,num_pieces=4,d_min=0.8,num_agents=N,slack_mode=1,slack_cost=50000.0,slack_decay=0.1)
    cfg=M.default_config(max_scenarios=S,**base); oc=O.base_config(**base)
    h=M.Handle(cfg)
    st,tg=dense_scenarios(S,N,N,0.8,spread=1.15,vmax=1.0)
    h.step_batch_host(st,tg)
    t=time.time(); got=h.step_batch_host(st,tg); dt=time.time()-t
    its=h.last_iteration_counts(S*N*2) if hasattr(h,'last_iteration_counts') else None
    Sc=min(S,8)
    ref=O.step_batch(oc,st[:Sc],tg[:Sc],threads=O.hardware_threads(),prec=1)
    ok=ref['n_curves']>0
    print(f"N={N} S={S}: {dt*1e3:.2f} ms, {S*N*2/dt:.3g} QPs/s; status got {np.bincount(got['status'].ravel(),minlength=8)} ref {np.bincount(ref['status'].ravel(),minlength=8)}")
    eq=np.array_equal(got['status'][:Sc],ref['status'])
    ex=np.abs(got['ctrl_pts'][:Sc]-ref['ctrl_pts'])
    print("   status equal",eq," err_x max",ex[ok].max() if ok.any() else None, " obj rel", (np.abs(got['objective'][:Sc]-ref['objective'])/np.maximum(1,np.abs(ref['objective'])))[ok].max() if ok.any() else None)
    if its is not None: print("   admm iters", np.percentile(its,[50,90,100]))
    h.close()
