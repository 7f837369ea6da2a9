import sys, numpy as np
sys.path.insert(0,'/root/repo')
import torch, bench
from bop_toolkit_b200 import batch, synth, _lib
dev=torch.device('cuda',0)
models=bench.build_models(); ms=batch.ModelSet(dev)
for m in models: ms.add(m['obj_id'], m['pts'], m['faces'], m['syms'], m['diameter'])
P=64
R_e,t_e,R_g,t_g,obj=bench.build_pairs(P,seed=0)
pe,pg=batch.pack_poses(R_e,t_e),batch.pack_poses(R_g,t_g)
K=synth.LM_K
depth=np.full((P,480,640),1000.0,np.float32)
batch.vsd_batch(ms,obj,pe,pg,depth,np.arange(P,dtype=np.int32),np.tile(K.reshape(1,9),(P,1)),15.0,bench.TAUS,True,'step','bop19',gate=True)
tot_c=0; tot_cost=0
for pair in range(0,P,7):
  for pose in (0,1):
    xy=batch.debug_triangle_list(ms,P,480,640,pair,pose).astype(np.int64)
    if len(xy)==0: continue
    minx=xy[:,[0,2,4]].min(1); maxx=xy[:,[0,2,4]].max(1); miny=xy[:,[1,3,5]].min(1); maxy=xy[:,[1,3,5]].max(1)
    w=np.minimum(639,maxx>>8)-np.maximum(0,(minx+255)>>8)+1; h=np.minimum(479,maxy>>8)-np.maximum(0,(miny+255)>>8)+1
    key=np.minimum(h,15)*16+np.minimum(w,15)
    srt=bool(np.all(np.diff(key)<=0))
    k=len(w)//32*32; W_=w[:k].reshape(-1,32); H_=h[:k].reshape(-1,32)
    cost=sum(np.where(H_>r,W_,0).max(1).sum() for r in range(H_.max()))*32
    tot_c+=(W_*H_).sum(); tot_cost+=cost
    if pair<8: print(pair,pose,len(xy),'sorted desc' if srt else 'NOT SORTED', 'first',list(zip(h[:6],w[:6])),'last',list(zip(h[-3:],w[-3:])))
print('nested lane efficiency of the actual lists: %.3f'%(tot_c/tot_cost))
