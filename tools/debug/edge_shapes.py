import sys; sys.path.insert(0,'/root/repo')
import numpy as np, torch
from syn3r_b200 import synth, warp as w, trajectory
from oracle import cpu_oracle as orc
dev=torch.device('cuda')
up=lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
# tiny and odd shapes
for (S,T,H,W) in [(1,1,1,1),(1,1,2,3),(1,2,5,33),(2,1,9,257),(1,1,49,129)]:
    tr=synth.make_trajectory(S,T,H,W,seed=5)
    ps,pt=synth.all_pairs(S,T)
    Kinv,par=w.forward_pair_params(tr.pose_src,tr.pose_tgt,tr.K,None,ps,pt)
    o=w.forward_warp_batched(up(w.pack_rgbd(tr.rgb,tr.depth)),up(Kinv),up(par),up(ps),want_winner=True,want_u8=True,want_f32=True)
    torch.cuda.synchronize()
    for p,(s,t) in enumerate(zip(ps,pt)):
        r=orc.forward_warp_ext([tr.rgb[s]],[None],[tr.depth[s]],[tr.pose_src[s]],tr.pose_tgt[t],[tr.K],None)
        assert np.array_equal(o['mask'][p].cpu().numpy().astype(bool), r['mask']), (S,T,H,W,p)
        assert np.max(np.abs(o['rgb_f32'][p].cpu().numpy()-r['out']),initial=0)<=255e-5
    iw=w.inverse_warp_batched(None,up(tr.depth),up(tr.depth_tgt),tr.K.astype(np.float32),up(w.inverse_pair_params(tr.pose_src,tr.pose_tgt,ps,pt)),up(ps),up(pt),want=("mask_reproj","soft_mask_reproj"))
    torch.cuda.synchronize()
    print('ok',(S,T,H,W))
# zero pairs
tr=synth.make_trajectory(1,1,8,8,seed=1)
try:
    e=np.zeros((0,),np.int32)
    Kinv,par=w.forward_pair_params(tr.pose_src,tr.pose_tgt,tr.K,None,e,e)
    o=w.forward_warp_batched(up(w.pack_rgbd(tr.rgb,tr.depth)),up(Kinv),up(par.reshape(0,24)),up(e),want_winner=True)
    torch.cuda.synchronize(); print('P=0 ok', {k:tuple(v.shape) for k,v in o.items()})
except Exception as ex:
    print('P=0 ->', type(ex).__name__, ex)
