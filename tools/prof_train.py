"""Self-training kernel alone: timing (CUDA events) and a short run for ncu.
   python tools/prof_train.py [log2 nets] [epochs] [reps]       # SRNN_LIB selects a library variant"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "self-replicating-neural-networks_b200"))
import torch  # noqa: E402
import srnn_engine as eng  # noqa: E402

lg = int(sys.argv[1]) if len(sys.argv) > 1 else 20
epochs = int(sys.argv[2]) if len(sys.argv) > 2 else 100
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
n = 1 << lg
probe = max(eng.fp32_probe(variant=0), eng.fp32_probe(variant=1))
KINDS = ((eng.KIND_WW, "WW(2,2)", 0, 1148.0), (eng.KIND_AGG, "AGG(4,2,2)", 4, 210.0), (eng.KIND_RNN, "RNN(2,2)", 0, None))
if os.environ.get("PROF_ONLY_WW"):
    KINDS = KINDS[:1]
for kind, name, agg, flop in KINDS:
    d = eng.make_desc(kind, 2, 2, agg)
    pop = eng.Population(d, n).init(seed=5)
    pop.train(2, perm_mode=eng.PERM_PHILOX, seed=2, want_loss=False)
    best = 1e9
    for _ in range(reps):
        pop.init(seed=5)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        pop.train(epochs, perm_mode=eng.PERM_PHILOX, seed=2, want_loss=False)
        e1.record()
        e1.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e-3)
    rate = n * epochs / best
    extra = "" if flop is None else "  %.1f TFLOP/s algorithmic = %.1f %% of the FFMA probe (%.1f TFLOP/s)" % (
        rate * flop / 1e12, 100 * rate * flop / (2 * probe), 2 * probe / 1e12)
    print("%-11s 2^%d nets x %d epochs: %.3f ms  %.3e epochs/s%s" % (name, lg, epochs, best * 1e3, rate, extra), flush=True)
