"""CPU tests that pin what CAN be pinned of the recurrent path -- and record what cannot.

The reference holds no RecurrentNeuralNetwork trajectory (every RNN block of code/setups/network_trajectorys.py is
under `if False`), only three counters: 46 divergent / 4 fix_zero after 100 self-attacks
(code/results/exp-applying_fixpoint-_1552681870.3570378-0/log.txt:9-10), 38 divergent / 12 other after 1000 train()
epochs (code/results/exp-training_fixpoint-_1552658296.0913951-0/log.txt:9-10), and the near-zero fixpoint fractions of
mixed-self-fixpoints (code/results/exp-mixed-self-fixpoints-_1552666977.5858653-0/log.txt:10).

1. The hand-written BPTT of the oracle (and, bit-for-bit, of the CUDA kernel) is the exact gradient of the loss Keras
   documents for `compile(loss='mse')` on a (1, T, 1) target -- mean over time of the squared error
   (code/network.py:566-574, 613-618) -- checked against an independent differentiator (torch autograd, float64).
2. The statistics these semantics produce are stable numbers, asserted here at n = 4000; the applying-fixpoints
   counter agrees with the reference; the training counter DOES NOT (0.49 divergent vs 38/50 = 0.76, 3.8 sigma of the
   reference's own sampling noise) and no variant of the semantics that Keras' documentation allows closes the gap
   (DESIGN.md section 5, table "RNN variants").  That disagreement is kept as a strict xfail so it cannot be
   forgotten: RNN *training* parity is UNPINNED.
"""
import numpy as np
import pytest

import oracle

T = 17


def _torch_forward(w, seq):
    """Stacked linear SimpleRNN(return_sequences=True), h_0 = 0: layers [K0 1x2, U0 2x2, K1 2x2, U1 2x2, K2 2x1, U2 1x1]
    in Keras get_weights() order (code/network.py:531-535)."""
    import torch
    n = w.shape[0]
    o, x = 0, seq[:, :, None]
    for i, u in ((1, 2), (2, 2), (2, 1)):
        K = w[:, o:o + i * u].reshape(n, i, u); o += i * u
        U = w[:, o:o + u * u].reshape(n, u, u); o += u * u
        h = torch.zeros(n, 1, u, dtype=w.dtype)
        outs = []
        for t in range(seq.shape[1]):
            h = torch.bmm(x[:, t:t + 1, :], K) + torch.bmm(h, U)
            outs.append(h)
        x = torch.cat(outs, 1)
    return x[:, :, 0]


def test_rnn_forward_matches_independent_restatement():
    import torch
    d = oracle.rnn()
    w = np.zeros((512, T), np.float32)
    oracle.init(d, w, seed=5)
    got = w.copy()
    oracle.self_apply(d, got, 1)
    want = _torch_forward(torch.from_numpy(w).double(), torch.from_numpy(w).double()).numpy()
    # running sums cancel, so the error is measured against the largest value of the net's own output sequence
    assert (np.abs(got - want) / np.abs(want).max(1, keepdims=True)).max() < 1e-5


def test_rnn_bptt_is_the_gradient_of_the_mean_squared_error_over_time():
    import torch
    d = oracle.rnn()
    n = 512
    w = np.zeros((n, T), np.float32)
    oracle.init(d, w, seed=6)
    w *= 0.5                       # keep first-step gradients moderate so f32 rounding stays far below the tolerance
    after = w.copy()
    loss = oracle.train(d, after, 1, lr=0.01)
    tw = torch.from_numpy(w).double().requires_grad_(True)
    y = torch.from_numpy(w).double()
    out = _torch_forward(tw, y)
    per_net = ((out - y) ** 2).mean(1)               # Keras: mean over the last axis, then over time and batch
    g, = torch.autograd.grad(per_net.sum(), tw)
    want = w.astype(np.float64) - np.float64(np.float32(0.01)) * g.numpy()
    step = np.abs(want - w).max(1, keepdims=True) + 1e-12
    assert (np.abs(after - want) / (np.abs(want) + step)).max() < 1e-5
    assert np.allclose(loss, per_net.detach().numpy(), rtol=1e-5)


def _fractions(n=4000, seed=11):
    d = oracle.rnn()
    w = np.zeros((n, T), np.float32)
    oracle.init(d, w, seed=seed)
    a = w.copy()
    oracle.self_apply(d, a, 100)
    _, ca = oracle.classify(d, a, 1e-4)
    b = w.copy()
    oracle.train(d, b, 1000)
    _, cb = oracle.classify(d, b, 1e-4)
    return ca / n, cb / n


def test_rnn_statistics_of_the_documented_semantics():
    applied, trained = _fractions()
    # applying-fixpoints: reference 46 divergent / 4 fix_zero of 50
    assert abs(applied[0] - 0.92) < 3 * np.sqrt(0.92 * 0.08 / 50)
    assert applied[0] + applied[1] > 0.995
    # training-fixpoints under mean-over-time MSE, lr 0.01: the blow-up happens in the first <= 10 epochs and is a
    # property of the initial gradient's size; 0.495 +- 0.01 at n = 4000 (torch-autograd restatement: 0.496)
    assert abs(trained[0] - 0.495) < 0.035, trained
    assert trained[1] + trained[2] + trained[3] < 0.01      # no fixpoints after 1000 epochs (reference: 0 of 50)


@pytest.mark.xfail(strict=True, reason="RNN training parity UNPINNED: reference logged 38/50 divergent, Keras-documented "
                                       "semantics give 0.49 (DESIGN.md section 5)")
def test_rnn_training_statistic_matches_reference_log():
    _, trained = _fractions()
    p = trained[0]
    assert abs(38 / 50.0 - p) < 3 * np.sqrt(p * (1 - p) / 50)
