"""The HBM-bound regime: ONE step per trip through memory for every per-net kernel and kind, with
algorithmic bytes / time against the copy peak.  python tools/prof_single_step.py [log2 nets]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "self-replicating-neural-networks_b200"))
import torch  # noqa: E402
import srnn_engine as eng  # noqa: E402

lg = int(sys.argv[1]) if len(sys.argv) > 1 else 24
n = 1 << lg
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    PEAK = 6650.0


def timed(fn, reps=10):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / reps


def line(name, what, t, nbytes):
    print("%-11s %-34s %7.3f ms  %.3e /s  %5.0f GB/s = %4.1f %% of %.0f" % (name, what, t * 1e3, n / t, nbytes / t / 1e9,
                                                                         100 * nbytes / t / 1e9 / PEAK, PEAK))


for name, desc in (("WW(2,2)", eng.make_desc(eng.KIND_WW, 2, 2)), ("AGG(4,2,2)", eng.make_desc(eng.KIND_AGG, 2, 2, 4)),
                   ("RNN(2,2)", eng.make_desc(eng.KIND_RNN, 2, 2))):
    nb = desc.n_weights * 4
    pop = eng.Population(desc, n).init(seed=1)
    tch = eng.Population(desc, n).init(seed=2)
    line(name, "init (write)", timed(lambda: pop.init(seed=1)), n * nb)
    pop.w.mul_(0.5)   # keep the iterates finite over the repetitions
    line(name, "self_apply x1 (read+write)", timed(lambda: pop.self_apply(1)), 2 * n * nb)
    pop.init(seed=1)
    line(name, "classify (read)", timed(lambda: pop.classify(1e-4)), n * nb)
    line(name, "train 1 epoch (read+write)", timed(lambda: pop.train(1, perm_mode=eng.PERM_PHILOX, seed=3, want_loss=False)), 2 * n * nb)
    line(name, "learn_from 1 epoch, arena order", timed(lambda: pop.train(1, perm_mode=eng.PERM_PHILOX, seed=3, teacher=tch, want_loss=False)),
         3 * n * nb)
    idx = torch.randperm(n, device="cuda", dtype=torch.int64).to(torch.int32)
    line(name, "learn_from 1 epoch, random teacher", timed(lambda: pop.train(1, perm_mode=eng.PERM_PHILOX, seed=3, teacher=tch, teacher_idx=idx,
                                                                             want_loss=False)), 3 * n * nb)
    del pop, tch, idx
