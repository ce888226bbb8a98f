"""Small driver for profiling the soup kernels: python tools/prof_soup.py [particles] [generations] [train]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "self-replicating-neural-networks_b200"))
import torch  # noqa: E402
import srnn_engine as eng  # noqa: E402
from soup import PopulationSoup  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
gens = int(sys.argv[2]) if len(sys.argv) > 2 else 6
train = int(sys.argv[3]) if len(sys.argv) > 3 else 0
path = int(sys.argv[4]) if len(sys.argv) > 4 else 0   # 0 auto, 1 fused kernel, 2 multi-pass, 3 persistent
order = sys.argv[5] if len(sys.argv) > 5 else "snapshot"   # or "exact"
ar = float(os.environ.get("PROF_ATTACK", "0.1"))   # rates overridable for phase-cost breakdowns
lfr = float(os.environ.get("PROF_LEARN", "0.1"))
kind = int(os.environ.get("PROF_KIND", "0"))       # 0 WW(2,2), 1 AGG(4,2,2), 2 RNN(2,2)
s = PopulationSoup(kind, n, aggregates=4 if kind == 1 else 0, seed=4, attacking_rate=ar, learn_from_rate=lfr, train=train, remove_divergent=True,
                   remove_zero=True, epsilon=1e-4, engine_path=path, order=order).seed_particles()
s.evolve(2)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
s.evolve(gens)
e1.record()
e1.synchronize()
print("kind %d attack %g learn %g" % (kind, ar, lfr), "particles %d train %d path %d order %s: %.3f ms / generation (rounds %d)" % (n, train, path, order,
      e0.elapsed_time(e1) / gens, s.rounds), s.count())
