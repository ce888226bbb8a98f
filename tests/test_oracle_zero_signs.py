"""CPU test of the lemma behind the fixed-length kernel's short inner steps (`ww_apply` ANYZERO):
when does the EXACT weightwise self-application not see the signs of the zeros it is given?

The kernel runs steps 1 .. n-1 of a fixed-length chain in a form that may leave a zero with either sign, and step n in
the exact form.  That is bit-identical to the all-exact chain if (A) the sign of a zero never changes a non-zero value, a
NaN / Inf or the zero / non-zero pattern of the next state, and (B) the exact form's own zeros do not depend on the
signs of the zeros in its input.  (A) holds for every shape.  (B) holds when the single output unit has at most two
inputs -- out = (+0 + p0) + p1 and +0 + p0 is never -0 -- and FAILS with three: a partial sum can underflow to -0 and the
third product can be a zero of either sign.  The oracle is the exact form; this test flips zero signs at random in
adversarial states (many zeros, underflowing products, Inf / NaN rows) and compares successors: magnitudes for every
shape, every bit for the width-2 shapes, and it shows the width-3 counterexamples -- which is why
`exact_step_ignores_zero_signs` (csrc/srnn_device.cuh) only admits shapes whose last layer has <= 2 inputs.
(The GPU side -- chains of the kernel match the oracle -- is tests/test_gpu_parity.py::test_self_apply_fixed_zero_signs.)"""
import numpy as np
import pytest

import oracle
from util_parity import assert_same_weights, glorot_population


@pytest.mark.parametrize("key", [(0, 2, 2, 0), (0, 3, 2, 0), (0, 2, 3, 0), (0, 2, 1, 0)])
def test_exact_step_and_the_signs_of_zeros(key):
    d = oracle.desc(*key)
    rng = np.random.default_rng(11)
    n, nw = 20000, d.n_weights
    two_inputs = key[1] <= 2                     # the output unit's fan-in is the width
    w = glorot_population(key, n, seed=12)
    w = (w * (np.float32(10.0) ** rng.integers(-24, 1, size=(n, 1)).astype(np.float32))).astype(np.float32)
    w[rng.random((n, nw)) < 0.2] = np.float32(0.0)
    w[5, 0], w[6, 1], w[7, nw - 1] = np.inf, np.nan, -np.inf
    state = w
    flipped, sign_only_differences = 0, 0
    for step in range(6):
        nxt = state.copy()
        oracle.self_apply(d, nxt, 1)
        twin = state.copy()
        flip = (twin == 0) & (rng.random(twin.shape) < 0.5)
        twin[flip] = -twin[flip]                      # +0 <-> -0
        flipped += int(flip.sum())
        oracle.self_apply(d, twin, 1)
        # (A): same NaN pattern, same magnitudes everywhere
        assert np.array_equal(np.isnan(twin), np.isnan(nxt))
        ok = ~np.isnan(nxt)
        assert np.array_equal(np.abs(twin)[ok].view(np.uint32), np.abs(nxt)[ok].view(np.uint32)), "%r step %d" % (key, step)
        if two_inputs:
            assert_same_weights(twin, nxt, "%r step %d" % (key, step))      # (B): every bit
        else:
            sign_only_differences += int((twin.view(np.uint32)[ok] != nxt.view(np.uint32)[ok]).sum())
        state = nxt
    assert flipped > 1000
    if not two_inputs:
        assert sign_only_differences > 0, "the width-3 counterexamples to (B) should show up in this population"


@pytest.mark.parametrize("key", [(0, 2, 2, 0), (0, 2, 3, 0), (0, 2, 1, 0), (0, 3, 2, 0)])
def test_short_inner_steps_then_one_exact_step(key):
    """The kernels' schedule itself, emulated on the CPU (oracle.self_apply_short_form): steps 1 .. n-1 with the output
    sum's first add spelled fma(x0, W0, +0), step n exact.  Width-2 shapes: equal to n exact steps in every bit, on 300 000
    adversarial nets and 2 .. 12 steps -- although the short form alone does leave zeros with the other sign (shown by
    running it without the exact finish).  WW(3,2): the schedule is NOT admitted (exact_step_ignores_zero_signs); here it
    only has to keep magnitudes."""
    d = oracle.desc(*key)
    rng = np.random.default_rng(21)
    n, nw = 300000, d.n_weights
    w = glorot_population(key, n, seed=22)
    w = (w * (np.float32(10.0) ** rng.integers(-26, 1, size=(n, 1)).astype(np.float32))).astype(np.float32)
    w[rng.random((n, nw)) < 0.1] = np.float32(-0.0)
    w[rng.random((n, nw)) < 0.1] = np.float32(0.0)
    w[5, 0], w[6, 1], w[7, nw - 1] = np.inf, np.nan, -np.inf
    sign_differences_without_the_finish = 0
    for steps in (2, 3, 5, 12):
        exact, short, unfinished = w.copy(), w.copy(), w.copy()
        oracle.self_apply(d, exact, steps)
        oracle.self_apply_short_form(d, short, steps)
        oracle.self_apply_short_form(d, unfinished, steps, exact_last=False)
        ok = ~np.isnan(exact)
        for other in (short, unfinished):                       # (A): magnitudes and NaNs never move
            assert np.array_equal(np.isnan(other), np.isnan(exact))
            assert np.array_equal(np.abs(other)[ok].view(np.uint32), np.abs(exact)[ok].view(np.uint32))
        sign_differences_without_the_finish += int((unfinished.view(np.uint32)[ok] != exact.view(np.uint32)[ok]).sum())
        if key[1] <= 2:
            assert_same_weights(short, exact, "%r, %d steps" % (key, steps))
    assert sign_differences_without_the_finish > 0, "the population is meant to make the short form show"
