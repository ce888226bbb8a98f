"""GPU tests: the reference's experiment scripts (code/setups/*.py) in miniature, written against the drop-in
class API exactly the way the scripts use it, and checked against the results the reference logged
(tests/golden/logs.json) within sampling noise."""
import json
import os
import random

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api():
    import network
    import soup
    import experiment
    network.set_seed(4321)
    random.seed(7)
    np.random.seed(7)
    return network, soup, experiment


@pytest.fixture(scope="module")
def logs(golden_dir=None):
    here = os.path.dirname(os.path.abspath(__file__))
    return json.load(open(os.path.join(here, "golden", "logs.json")))


def generate_counters():
    return {'divergent': 0, 'fix_zero': 0, 'fix_other': 0, 'fix_sec': 0, 'other': 0}


def count(counters, net, notable_nets=[]):
    """The count() every setup defines (e.g. code/setups/applying-fixpoints.py:16-30)."""
    if net.is_diverged():
        counters['divergent'] += 1
    elif net.is_fixpoint():
        if net.is_zero():
            counters['fix_zero'] += 1
        else:
            counters['fix_other'] += 1
    elif net.is_fixpoint(2):
        counters['fix_sec'] += 1
    else:
        counters['other'] += 1
    return counters, notable_nets


def _generators(network):
    return [lambda: network.WeightwiseNeuralNetwork(width=2, depth=2).with_keras_params(activation='linear', use_bias=False),
            lambda: network.AggregatingNeuralNetwork(aggregates=4, width=2, depth=2).with_keras_params(activation='linear', use_bias=False),
            lambda: network.RecurrentNeuralNetwork(width=2, depth=2).with_keras_params(activation='linear', use_bias=False)]


def test_applying_fixpoints(api, logs):
    """applying-fixpoints.py: trials x 100 self_attack, then count (eps 1e-4).  Reference, 50 trials:
    WW 23/27, AGG 4/46, RNN 46/4 (divergent / fix_zero)."""
    network, _, _ = api
    ref = logs["applying_fixpoint"]["counters"]
    trials = 150
    for gen in _generators(network):
        counters = generate_counters()
        for _ in range(trials):
            net = network.ParticleDecorator(gen())
            net.with_params(epsilon=1e-4)
            name = net.net.__class__.__name__
            for _ in range(100):
                net.self_attack()
            count(counters, net)
        network.K.clear_session()
        p_ref = ref[name]["divergent"] / 50.0
        p = counters["divergent"] / trials
        sigma = np.sqrt(max(p_ref * (1 - p_ref), 0.02) * (1 / 50.0 + 1 / trials))
        assert abs(p - p_ref) < 4 * sigma, (name, counters, ref[name])
        assert counters["divergent"] + counters["fix_zero"] >= trials - 3


def test_training_fixpoints(api, logs):
    """training-fixpoints.py: 1000 x train(), then count: WW -> fix_other (50/50), AGG -> other (50/50)."""
    network, _, _ = api
    for gen, want in zip(_generators(network)[:2], ("fix_other", "other")):
        counters = generate_counters()
        for _ in range(6):
            net = network.TrainingNeuralNetworkDecorator(network.ParticleDecorator(gen())).with_params(epsilon=1e-4)
            for run_id in range(1000):
                loss = net.compiled().train(epoch=run_id + 1)
            count(counters, net)
        assert counters[want] == 6, counters
        assert np.isfinite(loss)


def test_mixed_self_fixpoints(api, logs):
    """mixed-self-fixpoints.py: 4 rounds of {self_attack, t x train}: more training -> more fixpoints
    (reference WW: 0.2 of the nets at t = 0, >= 0.7 from t = 200)."""
    network, _, _ = api
    gen = _generators(network)[0]
    frac = {}
    for t in (0, 200):
        counters = generate_counters()
        trials = 12
        for _ in range(trials):
            net = network.TrainingNeuralNetworkDecorator(network.ParticleDecorator(gen())).with_params(epsilon=1e-4)
            for selfattack_id in range(4):
                net.self_attack()
                for train_id in range(t):
                    net.compiled().train(epoch=selfattack_id * t + train_id)
                if net.is_diverged() or net.is_fixpoint():
                    break
            count(counters, net)
        frac[t] = (counters['fix_zero'] + counters['fix_other']) / trials
    assert frac[200] >= 0.5 and frac[200] >= frac[0], frac


def test_learn_from_soup_severity(api, logs):
    """learn_from_soup.py: soup of 10, 100 generations, learn_from_rate 0.1: with severity 0 no non-zero fixpoints
    (reference 0.0), with severity 50 almost all particles (reference 9.1 of 10)."""
    network, soup_mod, _ = api
    res = {}
    for severity in (0, 50):
        nonzero = 0
        for trial in range(3):
            gen = lambda: network.TrainingNeuralNetworkDecorator(network.WeightwiseNeuralNetwork(2, 2)) \
                .with_keras_params(activation='linear').with_params(epsilon=1e-4)  # noqa: E731
            s = soup_mod.Soup(10, gen).with_params(attacking_rate=-1, learn_from_rate=0.1, train=0,
                                                   learn_from_severity=severity)
            s.seed()
            for _ in range(100):
                s.evolve()
            for p in s.particles:
                if not p.is_diverged() and p.is_fixpoint() and not p.is_zero():
                    nonzero += 1
        res[severity] = nonzero / 3.0
    assert res[0] <= 1.0 and res[50] >= 6.0, res


def test_mixed_soup_trains(api, logs):
    """mixed-soup.py: soup of 10, 5 generations, attack 0.1, train t per generation: non-zero fixpoints grow with t
    (reference WW: 0 at t = 0, 8.8 of 10 at t = 100)."""
    network, soup_mod, _ = api
    res = {}
    for t in (0, 100):
        total = 0
        for trial in range(3):
            gen = lambda: network.TrainingNeuralNetworkDecorator(network.WeightwiseNeuralNetwork(2, 2)) \
                .with_params(epsilon=1e-4)  # noqa: E731
            s = soup_mod.Soup(10, gen)
            s.with_params(attacking_rate=0.1, learn_from_rate=-1, train=t, learn_from_severity=-1)
            s.seed()
            for _ in range(5):
                s.evolve()
            for net in s.particles:
                if not net.is_diverged() and net.is_fixpoint() and not net.is_zero():
                    total += 1
        res[t] = total / 3.0
    assert res[0] <= 1.0 and res[100] >= 5.0, res


def test_fixpoint_density_fresh_nets(api):
    """fixpoint-density.py: classify fresh nets without any self-application (eps 1e-4): essentially all 'other'."""
    network, _, _ = api
    for gen in _generators(network)[:2]:
        counters = generate_counters()
        for _ in range(300):
            net = network.ParticleDecorator(gen().with_params(epsilon=1e-4))
            count(counters, net)
            network.K.clear_session()
        assert counters['other'] >= 295, counters
