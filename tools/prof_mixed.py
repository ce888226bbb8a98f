"""Mixed 10 M soup (WW / AGG / RNN thirds) on one GPU: python tools/prof_mixed.py [particles per kind] [generations]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "self-replicating-neural-networks_b200"))
import torch  # noqa: E402
from soup import PopulationSoup, evolve_together  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 3333328
gens = int(sys.argv[2]) if len(sys.argv) > 2 else 20
kw = dict(attacking_rate=0.1, learn_from_rate=0.1, train=0, learn_from_severity=1, remove_divergent=True, remove_zero=True,
          epsilon=1e-4)
soups = [PopulationSoup(k, n, aggregates=4 if k == 1 else 0, seed=4, **kw).seed_particles() for k in (0, 1, 2)]


def timed(fn):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1)


evolve_together(soups, 2)
t = min(timed(lambda: evolve_together(soups, gens)) for _ in range(3))
print("together: %.3f ms / generation (%.0f generations/s)" % (t / gens, 1e3 * gens / t))
for s in soups:
    s.evolve(2)
t = min(timed(lambda: [s.evolve(gens) for s in soups]) for _ in range(3))
print("one kind after the other: %.3f ms / generation (%.0f generations/s)" % (t / gens, 1e3 * gens / t))
