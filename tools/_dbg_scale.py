import sys, torch, numpy as np
sys.path.insert(0, '/root/repo')
from grid_b200 import engine
from oracle import cohort_check as C
SEED=20251104
n,b=3202,400000
dev=torch.device('cuda')
depth=engine.synth_depth(SEED,n,b,0,0,device=dev)
p=engine.Params()
res=engine.normalize(depth,b,p,want_zq=False,want_zh=False)
chk=C.CohortCheck(lambda r0,nr,c0,nc: engine.synth_depth(SEED,nr,nc,r0,c0,device=dev), n,b,dict(num_neighbors=10,n_nbr=5),dev)
out=chk.run([0,5])
sq=res.scale_q.cpu().numpy(); w=out['scale_q']
bad=np.nonzero(sq!=w)[0]
print('mismatch',len(bad), 'ties', len(out['tie_rows']))
rs=res.extra['rowsum'].cpu().numpy(); rc=res.extra['rowcnt'].cpu().numpy()
for i in bad[:8]:
    print(i, repr(sq[i]), repr(w[i]), rs[i], rc[i], repr(rs[i]/(100.0*rc[i])), repr((rs[i]/100.0)/rc[i]), repr(res.raw_mean[i].item()))
