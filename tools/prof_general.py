"""Timing of the general self-apply kernel (run_net rule) and the train kernel for the three kinds: 2^22 nets."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "self-replicating-neural-networks_b200"))
import torch  # noqa: E402
import srnn_engine as eng  # noqa: E402

n = 1 << 22


def timed(prep, fn, reps=5):
    tot = 0.0
    for _ in range(reps + 1):
        prep()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        tot += e0.elapsed_time(e1) if _ else 0.0
    return tot / reps


for name, d in (("WW", eng.make_desc(0, 2, 2)), ("AGG", eng.make_desc(1, 2, 2, 4)), ("RNN", eng.make_desc(2, 2, 2))):
    pop = eng.Population(d, n)
    t1 = timed(lambda: pop.init(seed=2), lambda: pop.self_apply(20, mode=eng.RUN_UNTIL, eps=1e-4, want_steps=True, want_counters=True))
    t2 = timed(lambda: pop.init(seed=2), lambda: pop.self_apply(100, mode=eng.RUN_UNTIL, eps=1e-4, want_counters=True))
    t3 = timed(lambda: pop.init(seed=3), lambda: pop.train(10, perm_mode=eng.PERM_PHILOX, seed=4, want_loss=True))
    t4 = timed(lambda: pop.init(seed=3), lambda: pop.classify(1e-4))
    print("%-4s general(20, until) %.3f ms   general(100, until) %.3f ms   train(10) %.3f ms   classify %.3f ms" % (name, t1, t2, t3, t4))
