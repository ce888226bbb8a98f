"""One launch of every hot-path kernel at population scale, for ncu surveys:
   ncu --set full -k regex:srnn python tools/prof_kernels.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "self-replicating-neural-networks_b200"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import srnn_engine as eng  # noqa: E402
from soup import PopulationSoup  # noqa: E402

n = 1 << 22
dw = eng.make_desc(eng.KIND_WW, 2, 2)
da = eng.make_desc(eng.KIND_AGG, 2, 2, 4)
dr = eng.make_desc(eng.KIND_RNN, 2, 2)
for d in (dw, da, dr):
    pop = eng.Population(d, n).init(seed=1)                        # init_kernel
    pop.self_apply(20)                                             # self_apply_fixed_kernel
    pop.init(seed=2)
    pop.self_apply(20, mode=eng.RUN_UNTIL, eps=1e-4, want_steps=True, want_counters=True)   # general kernel
    pop.init(seed=3)
    pop.classify(1e-4)                                             # classify_kernel
    pop.train(10, perm_mode=eng.PERM_PHILOX, seed=4, want_loss=True)   # train_kernel
    idx = torch.randperm(n, device="cuda", dtype=torch.int64).to(torch.int32)
    half = n // 2
    pop.attack(idx[:half], idx[half:])                             # apply_kernel (random pairs)
    del pop
base = np.zeros(14, np.float32)
base[[0, 8, 12]] = 1
eng.kfv(dw, base, n, 1e-3, 3, 100, 1e-4, seed=5)                  # kfv_kernel
s = PopulationSoup(eng.KIND_WW, n, seed=6, attacking_rate=0.1, learn_from_rate=0.1, train=1, remove_divergent=True,
                   remove_zero=True, epsilon=1e-4).seed_particles()
s.evolve(3)                                                        # soup_lists_kernel + soup_sync_kernel
torch.cuda.synchronize()
print("done")
