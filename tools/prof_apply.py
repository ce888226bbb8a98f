"""Timing of the pair-apply kernel (srnn_apply = attack for a list of pairs): python tools/prof_apply.py [log2 nets]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "self-replicating-neural-networks_b200"))
import torch  # noqa: E402
import srnn_engine as eng  # noqa: E402

lg = int(sys.argv[1]) if len(sys.argv) > 1 else 24
n = 1 << lg


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


for name, desc in (("WW(2,2)", eng.make_desc(eng.KIND_WW, 2, 2)), ("AGG(4,2,2)", eng.make_desc(eng.KIND_AGG, 2, 2, 4)),
                   ("RNN(2,2)", eng.make_desc(eng.KIND_RNN, 2, 2))):
    pop = eng.Population(desc, n).init(seed=1)
    out = eng.Population(desc, n)
    idx = torch.randperm(n, device="cuda", dtype=torch.int64).to(torch.int32)
    half = n // 2
    nb = desc.n_weights * 4
    # (a) random pairs, in place (attack): three gathers / scatters of one record each per pair
    t = timed(lambda: pop.attack(idx[:half], idx[half:]))
    print("%-11s random pairs in place : %7.3f ms  %.3e pairs/s  %6.0f GB/s algorithmic (3 records per pair)"
          % (name, t * 1e3, half / t, half * 3 * nb / t / 1e9))
    # (b) random attackers, victims and results in arena order
    t = timed(lambda: pop.apply_pairs(idx, pop, None, out, None, n=n))
    print("%-11s random attacker       : %7.3f ms  %.3e pairs/s  %6.0f GB/s algorithmic" % (name, t * 1e3, n / t, n * 3 * nb / t / 1e9))
    # (c) everything in arena order (attacker i, victim i -> out i)
    t = timed(lambda: pop.apply_pairs(None, pop, None, out, None, n=n))
    print("%-11s arena order           : %7.3f ms  %.3e pairs/s  %6.0f GB/s algorithmic (2 records per pair)"
          % (name, t * 1e3, n / t, n * 2 * nb / t / 1e9))
    del pop, out
