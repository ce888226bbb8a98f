"""Reference-STRUCTURED CPU twin (test / baseline infrastructure, not the product, not the parity oracle).

Same per-object Python loops as the reference's WeightwiseNeuralNetwork.apply_to_weights
(code/network.py:239-279: one model.predict per weight point from a triple loop), with a float32 NumPy dot in
place of the Keras call.  It exists only to time "what the reference's control flow costs on a CPU core"
(BASELINE.md section 4, item 2); real Keras-TF1 adds 0.3-1 ms of session overhead per predict on top.
Results agree with the C oracle to ~1e-6 relative (NumPy's dot does not follow the pinned FMA placement).
"""
import time

import numpy as np

SHAPES = [(4, 2), (2, 2), (2, 1)]


def _norm(v, m):
    return float(v) / float(m) if m > 1 else float(v)


def apply_to_weights(own, old):
    new = [m.copy() for m in own]
    top_l = len(old) - 1
    for li, layer in enumerate(old):
        top_c = len(layer) - 1
        for ci, cell in enumerate(layer):
            top_w = len(cell) - 1
            for wi, weight in enumerate(cell):
                x = np.array([[weight, _norm(li, top_l), _norm(ci, top_c), _norm(wi, top_w)]], dtype=np.float32)
                for m in own:                       # "model.predict": 4 -> 2 -> 2 -> 1, linear, no bias
                    x = x.dot(m)
                new[li][ci][wi] = x[0][0]
    return new


def self_applications_per_second(n_nets=20, n_steps=20, seed=0):
    rng = np.random.default_rng(seed)
    nets = [[rng.uniform(-1, 1, s).astype(np.float32) for s in SHAPES] for _ in range(n_nets)]
    t0 = time.perf_counter()
    with np.errstate(all="ignore"):
        for w in nets:
            for _ in range(n_steps):
                w[:] = apply_to_weights(w, w)
    return n_nets * n_steps / (time.perf_counter() - t0)


if __name__ == "__main__":
    print("%.0f self-applications/s on one core (reference-structured NumPy twin)" % self_applications_per_second())
