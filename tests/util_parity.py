"""Shared comparison helpers for the parity tests."""
import numpy as np


def assert_same_weights(got, want, what=""):
    """Bit-exact parity for FP32 weights: identical bit patterns everywhere except that NaN matches NaN
    regardless of payload (x86 and the GPU produce different default-NaN payloads; the reference only
    ever tests np.isnan)."""
    got = np.ascontiguousarray(got, dtype=np.float32)
    want = np.ascontiguousarray(want, dtype=np.float32)
    assert got.shape == want.shape, (what, got.shape, want.shape)
    gn, wn = np.isnan(got), np.isnan(want)
    assert np.array_equal(gn, wn), "%s: NaN pattern differs at %d places" % (what, int((gn != wn).sum()))
    gb, wb = got.view(np.uint32)[~gn], want.view(np.uint32)[~wn]
    bad = gb != wb
    if bad.any():
        idx = np.flatnonzero(bad)[:5]
        raise AssertionError("%s: %d of %d values differ bitwise, e.g. got %r want %r" % (
            what, int(bad.sum()), gb.size, got[~gn][idx], want[~wn][idx]))


def glorot_population(desc_key, n, seed):
    """Random population with the Keras default init ranges (numpy RNG; only used as test input)."""
    kind, width, depth, aggs = desc_key
    rng = np.random.default_rng(seed)
    io = 4 if kind == 0 else (aggs if kind == 1 else 1)
    cols = []
    for l in range(depth + 1):
        fi = io if l == 0 else width
        fo = (1 if kind == 0 else io) if l == depth else width
        lim = np.sqrt(6.0 / (fi + fo))
        cols.append(rng.uniform(-lim, lim, (n, fi * fo)))
        if kind == 2:
            cols.append(rng.uniform(-1, 1, (n, fo * fo)))
    return np.ascontiguousarray(np.concatenate(cols, axis=1), dtype=np.float32)
