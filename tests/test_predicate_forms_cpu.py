"""CPU check of the predicate rewrite (csrc/srnn_device.cuh: is_diverged / is_zero / is_close as ONE max reduction and one
compare each, instead of a compare per weight): both forms, written out in NumPy with the kernels' flush-to-zero
compares, give the same booleans on adversarial vectors -- finite, +-0, denormal, +-Inf and NaN entries in every mix.

  m  = NaN-PROPAGATING max_k |w_k|  (PTX max.NaN.f32 -> np.maximum):  diverged <=> !(m <= FLT_MAX);  zero <=> m <= eps
  md = NaN-IGNORING    max_k |d_k|  (fmaxf           -> np.fmax), d = nw - w:  close <=> !diverged(nw) and !(md >= eps)

The GPU kernels are compared with the oracle's element-wise predicates in tests/test_gpu_parity.py; this file pins the
algebra the rewrite rests on, NaN corner cases included."""
import numpy as np

FLT_MAX = np.float32(3.402823466e+38)
FLT_MIN = np.float32(1.17549435e-38)


def _ftz(x):
    """What a flush-to-zero FP32 compare sees: denormals are zero."""
    x = x.copy()
    x[np.abs(x) < FLT_MIN] = 0
    return x


def _vectors(rng, n, k):
    scale = np.float32(10.0) ** rng.integers(-44, 39, size=(n, k)).astype(np.float32)
    v = (rng.standard_normal((n, k)).astype(np.float32) * scale).astype(np.float32)
    special = np.array([0.0, -0.0, np.inf, -np.inf, np.nan, 1e-45, -1e-45, 1e-4, -1e-4, 3.4e38], np.float32)
    pick = rng.random((n, k)) < 0.15
    v[pick] = special[rng.integers(0, len(special), size=int(pick.sum()))]
    v[rng.random(n) < 0.05] = np.nan                        # rows that are NaN everywhere
    return v


def test_one_number_predicates_equal_elementwise_predicates():
    rng = np.random.default_rng(1)
    n, k = 200000, 14
    with np.errstate(all="ignore"):
        w, nw = _vectors(rng, n, k), _vectors(rng, n, k)
        same = rng.random(n) < 0.3
        nw[same] = w[same] + (rng.standard_normal((int(same.sum()), k)) * 1e-5).astype(np.float32)   # near fixpoints
        for eps in (np.float32(1e-4), np.float32(1e-14), np.float32(0.5), FLT_MIN):
            aw, anw = np.abs(_ftz(w)), np.abs(_ftz(nw))
            # element-wise forms (the reference's: code/network.py:43-62, 133-157; NaN compares false)
            div_e = (~(aw <= FLT_MAX)).any(1)
            zero_e = (aw <= eps).all(1)
            d = np.abs(_ftz((nw - w).astype(np.float32)))
            close_e = ~(~(anw <= FLT_MAX)).any(1) & (~(d >= eps)).all(1)
            # one-number forms
            m = np.maximum.reduce(aw, axis=1)                 # NaN-propagating
            div_m = ~(m <= FLT_MAX)
            zero_m = m <= eps
            mn = np.maximum.reduce(anw, axis=1)
            md = np.fmax.reduce(d, axis=1)                     # NaN-ignoring
            close_m = (mn <= FLT_MAX) & ~(md >= eps)
            assert np.array_equal(div_e, div_m)
            assert np.array_equal(zero_e, zero_m)
            assert np.array_equal(close_e, close_m)
            assert div_e.any() and (~div_e).any() and zero_e.any() is not None and close_e.any()
