"""Every kernel once at tiny sizes (for compute-sanitizer runs):  compute-sanitizer --tool memcheck python tools/sanity_small.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "self-replicating-neural-networks_b200"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import srnn_engine as eng  # noqa: E402
from soup import PopulationSoup  # noqa: E402

SHAPES = [(0, 2, 2, 0), (0, 3, 2, 0), (1, 2, 2, 4), (1, 2, 2, 3), (2, 2, 2, 0), (2, 2, 1, 0),      # instantiated
          (0, 4, 3, 0), (1, 3, 3, 5), (2, 3, 2, 0)]                                                # generic path
n = 300
for key in SHAPES:
    d = eng.make_desc(key[0], key[1], key[2], key[3])
    fast = eng.shape_supported(d) == 1
    pop = eng.Population(d, n).init(seed=1)
    pop.init(seed=2, idx=np.array([0, 5, 299]))
    pop.self_apply(3)
    pop.init(seed=3)
    pop.self_apply(5, mode=eng.RUN_UNTIL, eps=1e-4, want_steps=True, want_cls=True, want_counters=True, want_traj=True)
    pop.init(seed=4)
    pop.classify(1e-4, want_flags=True)
    pop.train(2, perm_mode=eng.PERM_PHILOX, seed=5)
    ns = d.n_weights if key[0] == 0 else 1
    pop.train(1, perm_mode=eng.PERM_EXPLICIT, perm=np.arange(ns, dtype=np.uint8)[None])
    tea = eng.Population(d, n).init(seed=6)
    pop.train(1, teacher=tea, teacher_idx=np.arange(n)[::-1].copy())
    pop.attack(np.arange(0, 100), np.arange(100, 200))
    io = 4 if key[0] == 0 else (key[3] if key[0] == 1 else d.n_weights)
    pop.forward(np.zeros((7, io), np.float32), idx=np.arange(7))
    padded = eng.Population(d, 17, ld=d.n_weights + 3).init(seed=7)
    padded.self_apply(2)
    padded.classify(1e-4)
    if fast:
        sp = eng.make_soup_params(attacking_rate=0.3, learn_from_rate=0.3, train=1, learn_from_severity=1,
                                  remove_divergent=True, remove_zero=True, epsilon=1e-4, perm_mode=2)
        att = np.where(np.arange(n) % 3 == 0, (np.arange(n) * 7) % n, -1)
        lrn = np.where(np.arange(n) % 4 == 0, (np.arange(n) * 5) % n, -1)
        pop.soup_ordered(sp, att, lrn, seed=8, gen=1, want_logged=True, spares=tea.w[:n].contiguous())
        for path in (0, 2, 3):
            s = PopulationSoup(key[0], n, width=key[1], depth=key[2], aggregates=key[3], seed=9, attacking_rate=0.3,
                               learn_from_rate=0.3, train=1, remove_divergent=True, remove_zero=True, engine_path=path)
            s.seed_particles().evolve(1)
            s.evolve(3)
            s.track_uids = True
            s.evolve(1)
            s.count()
        sh = [pop.w[:150].clone(), pop.w[150:].clone()]
        out = torch.zeros_like(pop.w)
        scratch = torch.zeros(eng.soup_scratch_len(n, 150), dtype=torch.int32, device="cuda")
        for r in range(2):
            eng.soup_sync_peers(d, sp, [t.data_ptr() for t in sh], 150, r, d.n_weights, out[r * 150:(r + 1) * 150], 3, 1, scratch)
base = np.zeros(14, np.float32)
base[[0, 8, 12]] = 1
eng.kfv(eng.make_desc(0, 2, 2), base, 500, 1e-2, 2, 100, 1e-4, seed=1, want_weights=True)
dead = (torch.rand(5000, device="cuda") < 0.3).to(torch.uint8)
uid = torch.zeros(5000, dtype=torch.int64, device="cuda")
tmp = torch.zeros(16, dtype=torch.int64, device="cuda")
eng.assign_uids(dead, uid, 100, tmp)
eng.assign_uids_static(dead, uid, 100)
eng.fp32_probe(iters=8, variant=0)
eng.fp32_probe(iters=8, variant=2)
torch.cuda.synchronize()
print("sanity run complete")
