"""Run-time-shape (generic) kernels: soup generation and known-fixpoint sweep throughput for shapes outside SRNN_SHAPES.
   python tools/prof_generic.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "self-replicating-neural-networks_b200"))
import torch  # noqa: E402
import srnn_engine as eng  # noqa: E402
from soup import PopulationSoup  # noqa: E402


def timed(fn):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) * 1e-3


for kind, width, depth, aggs, name in ((0, 3, 3, 0, "WW(3,3)"), (1, 3, 2, 6, "AGG(6,3,2)"), (2, 3, 2, 0, "RNN(3,2)"),
                                       (0, 2, 2, 0, "WW(2,2) [instantiated, for scale]")):
    d = eng.make_desc(kind, width, depth, aggs)
    n = 1_000_000
    s = PopulationSoup(kind, n, width=width, depth=depth, aggregates=aggs, seed=4, attacking_rate=0.1, learn_from_rate=0.1,
                       train=0, remove_divergent=True, remove_zero=True, epsilon=1e-4).seed_particles()
    s.evolve(2)
    t = min(timed(lambda: s.evolve(6)) for _ in range(3)) / 6
    pop = eng.Population(d, n).init(seed=3)
    pop.self_apply(2)
    ta = min(timed(lambda: pop.self_apply(10)) for _ in range(3))
    line = "%-34s %2d weights: soup generation of 1e6 particles %.3f ms; self_apply %.2e self-applications/s" % (
        name, d.n_weights, t * 1e3, n * 10 / ta)
    if kind == 0:
        base = np.zeros(d.n_weights, np.float32)
        off = 0
        for (r, c) in eng.layer_shapes(d):
            base[off] = 1.0
            off += r * c
        eng.kfv(d, base, 1_000_000, 1e-3, 3, 100, 1e-4, seed=1, want_times=False)
        tk = min(timed(lambda: eng.kfv(d, base, 4_000_000, 1e-3, 3, 100, 1e-4, seed=1, want_times=False)) for _ in range(3))
        line += "; kfv (noise 1e-3) %.2e trials/s" % (4_000_000 / tk)
    print(line, flush=True)
