import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from syn3r_b200 import synth, warp as w
S,T,H,W = [int(v) for v in sys.argv[1:5]]
seed=int(sys.argv[5]); kw=eval(sys.argv[6]) if len(sys.argv)>6 else {}
npairs=int(sys.argv[7]) if len(sys.argv)>7 else None
tr = synth.make_trajectory(S,T,H,W,seed=seed,pose_step=1.5)
ps,pt = synth.all_pairs(S,T)
if npairs: ps,pt=ps[:npairs],pt[:npairs]
dev=torch.device('cuda')
Kinv,par = w.forward_pair_params(tr.pose_src,tr.pose_tgt,tr.K,None,ps,pt)
up=lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
rgbd=up(w.pack_rgbd(tr.rgb,tr.depth))
o=w.forward_warp_batched(rgbd,up(Kinv),up(par),up(np.asarray(ps,np.int32)),want_u8=True,want_f32=True,**kw)
torch.cuda.synchronize()
print('ok', sys.argv[1:], float(o['mask'].float().mean()))
