"""Known-fixpoint-variation sweep and run_net-until-vergence alone (CUDA events): python tools/prof_kfv.py [trials per level]
   SRNN_LIB selects a library variant."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "self-replicating-neural-networks_b200"))
import torch  # noqa: E402
import srnn_engine as eng  # noqa: E402

nt = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000


def timed(fn):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) * 1e-3


d = eng.make_desc(eng.KIND_WW, 2, 2, 0)
base = np.zeros(14, np.float32)
base[[0, 8, 12]] = 1.0
eng.kfv(d, base, nt, 1.0, 0, 100, 1e-4, seed=11, want_times=False)
t, steps, scale = 0.0, 0.0, 1.0
for level in range(10):
    res = {}
    t += timed(lambda: res.update(eng.kfv(d, base, nt, scale, level, 100, 1e-4, seed=11, want_times=False)))
    h = res["hist"].cpu().numpy().astype(np.float64)
    steps += (h[0] * np.arange(101)).sum() + nt
    scale /= 10.0
print("kfv sweep: %.3e trials/s  %.3e executed self-applications/s" % (10 * nt / t, steps / t), flush=True)
n = 1 << 24
for eps in (1e-14, 1e-4):
    pop = eng.Population(d, n).init(seed=7)
    pop.self_apply(100, mode=eng.RUN_UNTIL, eps=eps, want_counters=True)
    best = 1e9
    for _ in range(3):
        pop.init(seed=7)
        best = min(best, timed(lambda: pop.self_apply(100, mode=eng.RUN_UNTIL, eps=eps, want_counters=True)))
    print("run_net until vergence, eps %g: %.3e nets/s" % (eps, n / best), flush=True)
pop = eng.Population(d, n).init(seed=7)
pop.classify(1e-4)
print("classify 2^24: %.3f ms" % (1e3 * min(timed(lambda: pop.classify(1e-4)) for _ in range(3))))
