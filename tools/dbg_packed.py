import sys, numpy as np
sys.path.insert(0, "self-replicating-neural-networks_b200"); sys.path.insert(0, "."); sys.path.insert(0, "tests")
import srnn_engine as eng, oracle
from util_parity import glorot_population
for key in [(0,2,2,0),(1,2,2,4),(2,2,2,0)]:
    dg=eng.make_desc(*key[:3], key[3]); do=oracle.desc(*key[:3], key[3])
    n=20000
    w=glorot_population(key,n,1)
    pop=eng.Population.from_numpy(dg,w)
    io = 4 if key[0]==0 else (4 if key[0]==1 else dg.n_weights)
    x=np.random.default_rng(0).uniform(-1,1,(n,io)).astype(np.float32)
    y=pop.forward(x).cpu().numpy(); yr=oracle.forward(do,w,x)
    print(key,'forward mismatches',(y.view(np.uint32)!=yr.view(np.uint32)).sum(),'of',y.size, np.abs(y-yr).max())
    pop.self_apply(1); ref=w.copy(); oracle.self_apply(do,ref,1)
    g=pop.numpy(); print('  self_apply(1) mismatches',(g.view(np.uint32)!=ref.view(np.uint32)).sum())
key=(0,2,2,0)
dg=eng.make_desc(*key[:3], key[3]); do=oracle.desc(*key[:3], key[3])
w=glorot_population(key,20000,1)
pop=eng.Population.from_numpy(dg,w); pop.self_apply(1); ref=w.copy(); oracle.self_apply(do,ref,1)
g=pop.numpy()
print('per-column mismatches', (g.view(np.uint32)!=ref.view(np.uint32)).sum(axis=0))
# same through apply_pairs and through forward with the exact codes
L=[0]*8+[.5]*4+[1,1]; R=[0,0,1/3,1/3,2/3,2/3,1,1,0,0,1,1,0,1]; C=[0,1]*6+[0,0]
for k in range(14):
    x=np.stack([w[:,k], np.full(20000,L[k],np.float32), np.full(20000,R[k],np.float32), np.full(20000,C[k],np.float32)],1).astype(np.float32)
    y=eng.Population.from_numpy(dg,w).forward(x).cpu().numpy()[:,0]
    print(k, 'forward-vs-oracle', (y.view(np.uint32)!=ref[:,k].view(np.uint32)).sum(), 'selfapply-vs-oracle', (g[:,k].view(np.uint32)!=ref[:,k].view(np.uint32)).sum())
