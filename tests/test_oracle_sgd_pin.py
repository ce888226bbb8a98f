"""CPU tests: the SGD arithmetic (DESIGN.md rule R7) pinned against the reference's own Keras
trajectories at -- and beyond -- the north_star's tolerance ("weights after N self-train steps match the reference
within 1e-5 relative").

The reference logged the weights of 10 WeightwiseNeuralNetwork(2,2) particles after each of 101 train() epochs
(code/results/self_training_weightwise_network/trajectorys.dill, produced by code/setups/network_trajectorys.py:53-67;
fixture tests/golden/ww_self_training.npz).  Keras shuffles the 14 samples of every epoch with the unseeded global NumPy
RNG, so the logs do not say in which order the samples ran.  tests/golden/make_sgd_perms.py recovered that order for
the first 20 epochs of every particle by an exhaustive meet-in-the-middle search over all 14! orders
(oracle/sgd_pin.c; fixture tests/golden/ww_sgd_perms.npz).  With the order known, one oracle epoch must land ON the
golden next state:

  * 185 of 199 epochs: bit for bit (all 14 weights, 0 ulp) -- which also decides the two roundings the forward goldens
    left open: the backward product g.W^T is a dot product (unfused), the update is W - rn(lr * dW) (two roundings);
    the alternatives are refuted on the same data (test_rounding_alternatives_are_refuted);
  * 193 of 199 within 1e-5 relative per weight;
  * 6 early epochs stay 1.5e-5 .. 1.2e-3 away with the best order found -- unexplained, listed, not hidden.
"""
import ctypes as C
import os
import sys

import numpy as np
import pytest

import oracle

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))


@pytest.fixture(scope="module")
def pins(golden_dir):
    g = np.load(os.path.join(golden_dir, "ww_self_training.npz"))
    p = np.load(os.path.join(golden_dir, "ww_sgd_perms.npz"))
    return g["weights"], p


def _oracle_epochs(W, p, rows):
    d = oracle.ww()
    out = np.empty((len(rows), 14), np.float32)
    for n, r in enumerate(rows):
        w = np.ascontiguousarray(W[p["before"][r]][None, :]).copy()
        oracle.train(d, w, 1, lr=0.01, perm_mode=oracle.PERM_EXPLICIT, perm=p["order"][r][None, :])
        out[n] = w[0]
    return out


def test_orders_are_permutations_of_consecutive_epochs(pins, golden_dir):
    W, p = pins
    g = np.load(os.path.join(golden_dir, "ww_self_training.npz"))
    assert len(p["before"]) >= 190
    assert (np.sort(p["order"], axis=1) == np.arange(14)).all()
    assert (g["particle"][p["before"]] == g["particle"][p["after"]]).all()
    assert (p["after"] == p["before"] + 1).all()


def test_oracle_epoch_reproduces_keras_bit_for_bit(pins):
    """The main C oracle (srnn_oracle_train, PERM_EXPLICIT) with the recovered order vs the golden next state."""
    W, p = pins
    rows = np.arange(len(p["before"]))
    got = _oracle_epochs(W, p, rows)
    want = W[p["after"]]
    exact = (got.view(np.uint32) == want.view(np.uint32)).all(axis=1)
    rel = (np.abs(got - want) / np.abs(want)).max(axis=1)
    assert exact.sum() >= 185, exact.sum()
    assert (rel <= 1e-5).sum() >= 193, np.sort(rel)[-10:]          # the north_star's tolerance, per weight
    assert rel.max() < 2e-3
    # the fixture's own bookkeeping (computed by oracle/sgd_pin.c, variant gp 1 / up 0) agrees with the main oracle
    assert np.array_equal(exact, p["ulp"][:, 2] == 0)
    # every particle contributes bit-exact epochs, and the very first (largest-gradient) epochs are among them
    first = [int(np.nonzero(p["before"] // 102 == u)[0][0]) for u in range(10)]
    assert exact[first].sum() >= 9


def test_rounding_alternatives_are_refuted(pins):
    """Same orders, other roundings: fused update, GEMM-style (fma) backward product.  Each loses dozens of epochs."""
    W, p = pins
    variants = [tuple(v) for v in p["variants"].tolist()]
    exact = {v: int((p["ulp"][:, i] == 0).sum()) for i, v in enumerate(variants)}
    assert exact[(1, 0)] >= 185
    assert exact[(0, 0)] <= exact[(1, 0)] - 20 and exact[(2, 0)] <= exact[(1, 0)] - 15      # fma in g.W^T: refuted
    assert max(exact[(0, 1)], exact[(1, 1)], exact[(2, 1)]) <= 100                           # fused update: refuted
    # and the library that produced the table still reproduces it
    import make_sgd_perms as M
    l = M.lib()
    for r in (0, 57, 140):
        a, b = W[p["before"][r]], W[p["after"][r]]
        row = [int(M.ulp_dist(M.epoch32(l, a, p["order"][r], gp=gp, up=up)[0], b).sum()) for gp, up in variants]
        assert row == p["ulp"][r].tolist()


def test_free_running_twenty_epochs_stay_within_tolerance(pins):
    """Chain the recovered orders: start from the golden state of epoch 0 and run the oracle alone for as long as the
    particle's epochs are bit-exact; the free-running state equals the golden one at every step (no drift at all)."""
    W, p = pins
    d = oracle.ww()
    longest = 0
    for u in range(10):
        rows = np.nonzero(p["before"] // 102 == u)[0]
        w = np.ascontiguousarray(W[p["before"][rows[0]]][None, :]).copy()
        run = 0
        for r in rows:
            if p["ulp"][r, 2] != 0 or p["before"][r] != p["before"][rows[0]] + run:
                break
            oracle.train(d, w, 1, lr=0.01, perm_mode=oracle.PERM_EXPLICIT, perm=p["order"][r][None, :])
            assert np.array_equal(w[0].view(np.uint32), W[p["after"][r]].view(np.uint32)), (u, run)
            run += 1
        longest = max(longest, run)
    assert longest >= 15
