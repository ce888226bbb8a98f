"""Self-application / self-training rates of the three benchmarked net kinds at population scale."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "self-replicating-neural-networks_b200"))
import torch
import srnn_engine as eng

def timed(fn):
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(); e1.record(); e1.synchronize(); return e0.elapsed_time(e1) * 1e-3

n = 1 << 23
for name, d in (("WW(2,2)", eng.make_desc(0, 2, 2)), ("AGG(4,2,2)", eng.make_desc(1, 2, 2, 4)), ("RNN(2,2)", eng.make_desc(2, 2, 2))):
    pop = eng.Population(d, n).init(seed=1)
    pop.self_apply(5)
    pop.init(seed=2)
    t = timed(lambda: pop.self_apply(100))
    pop.init(seed=3)
    pop.train(2, perm_mode=2, seed=1, want_loss=False)
    pop.init(seed=4)
    tt = timed(lambda: pop.train(20, perm_mode=2, seed=1, want_loss=False))
    print("%-11s self-applications/s %.3e   self-train epochs/s %.3e" % (name, n * 100 / t, n * 20 / tt))
