#!/usr/bin/env python
"""Extract golden vectors from the reference's committed experiment artifacts.

Run ONCE in the build container (where /root/reference exists); the resulting small
``*.npz`` / ``*.json`` files are committed and are what the tests read -- nothing at
test time touches /root/reference.

The reference (illiumst/self-replicating-neural-networks) has no test-suite; the only
things that pin its numerical behaviour are the ``*.dill`` trajectories and ``log.txt``
files its authors committed after real Keras runs.  This script
turns those into plain arrays:

  ww_self_application.npz   20 WW(2,2) particles, every saved state of run_net(100)
                            (code/setups/experiments/exp-weightwise_self_application-*/trajectorys.dill)
  agg_self_application.npz  10 AGG(4,2,2) particles, same
                            (code/results/self_application_aggregation_network/trajectorys.dill)
  ww_self_training.npz      10 WW(2,2) particles x 101 SGD epochs
                            (code/results/self_training_weightwise_network/trajectorys.dill)
  learn_from_soup.npz       soup of 10, 100 generations, learn_from events with counterpart
                            (code/results/exp-learn-from-soup-*/soup.dill)
  train_soup.npz            soup of 20, train=30, 100 generations, per-generation loss
                            (code/results/Soup/soup.dill)
  kfv.npz                   known-fixpoint-variation raw xs / ys / zs (1000 trials)
  logs.json                 every counter / curve the reference logged (behavioural KATs)

Usage:  python tests/golden/make_golden.py [/root/reference]
"""
import json
import os
import sys
import types

import numpy as np

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
CODE = os.path.join(REF, "code")
OUT = os.path.dirname(os.path.abspath(__file__))


def _install_stubs():
    """soup.dill pickles reference classes whose modules import TensorFlow (absent here).
    Register empty stand-ins so dill can resolve the names; code objects are not needed."""
    import dill

    class _Dummy:
        def __init__(self, *a, **k):
            pass

        def __setstate__(self, state):
            if isinstance(state, dict):
                self.__dict__.update(state)

    for name in ["tensorflow", "tensorflow.python", "tensorflow.python.keras",
                 "tensorflow.python.keras.models", "tensorflow.python.keras.callbacks",
                 "tensorflow.python.keras.layers", "tensorflow.python.keras.backend",
                 "keras", "keras.backend"]:
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["tensorflow.python.keras.models"].Sequential = _Dummy
    sys.modules["tensorflow.python.keras.callbacks"].Callback = _Dummy
    sys.modules["tensorflow.python.keras.layers"].Dense = _Dummy
    sys.modules["tensorflow.python.keras.layers"].SimpleRNN = _Dummy
    sys.modules["tensorflow.python.keras"].backend = sys.modules["tensorflow.python.keras.backend"]
    sys.modules["keras"].backend = sys.modules["keras.backend"]
    sys.modules["keras.backend"].clear_session = lambda: None

    # Python-3.6 code objects (15-field layout) cannot be rebuilt on 3.12: replace by placeholders.
    import dill._dill as dd
    orig_lt = dd._load_type

    def _lt(name):
        if name == "CodeType":
            return lambda *a, **k: (lambda *aa, **kk: None).__code__
        return orig_lt(name)
    dd._load_type = _lt
    orig_cf = dd._create_function

    def _cf(*a, **k):
        try:
            return orig_cf(*a, **k)
        except Exception:
            return lambda *aa, **kk: None
    dd._create_function = _cf
    return dill


def _load(dill, rel):
    with open(os.path.join(CODE, rel), "rb") as fh:
        return dill.load(fh)


def _pack_particles(hp):
    """historical_particles: uid -> list of state dicts  ==>  flat arrays."""
    ws, pid, tm, act, cp, loss = [], [], [], [], [], []
    for uid in sorted(hp):
        for st in hp[uid]:
            ws.append(np.asarray(st["weights"], dtype=np.float32))
            pid.append(uid)
            tm.append(st.get("time", -1) if st.get("time", -1) is not None else -1)
            act.append(str(st.get("action", "")))
            c = st.get("counterpart", None)
            cp.append(-1 if c is None else int(c))
            l = st.get("loss", None)
            loss.append(np.nan if l is None else float(l))
    return dict(weights=np.stack(ws), particle=np.asarray(pid, np.int64), time=np.asarray(tm, np.int64),
                action=np.asarray(act), counterpart=np.asarray(cp, np.int64),
                loss=np.asarray(loss, np.float64))


def main():
    sys.path.insert(0, CODE)
    dill = _install_stubs()
    logs = {}

    exp = _load(dill, "setups/experiments/exp-weightwise_self_application-_1552664922.4501734-0/trajectorys.dill")
    np.savez_compressed(os.path.join(OUT, "ww_self_application.npz"), **_pack_particles(exp.historical_particles))
    logs["ww_self_application"] = dict(counters=exp.counters, epsilon=1e-14, step_limit=100)

    exp = _load(dill, "results/self_application_aggregation_network/trajectorys.dill")
    np.savez_compressed(os.path.join(OUT, "agg_self_application.npz"), **_pack_particles(exp.historical_particles))
    logs["agg_self_application"] = dict(counters=exp.counters, epsilon=1e-14, step_limit=100)

    exp = _load(dill, "results/self_training_weightwise_network/trajectorys.dill")
    np.savez_compressed(os.path.join(OUT, "ww_self_training.npz"), **_pack_particles(exp.historical_particles))

    soup = _load(dill, "results/exp-learn-from-soup-_1552658566.5572753-0/soup.dill")
    np.savez_compressed(os.path.join(OUT, "learn_from_soup.npz"), **_pack_particles(soup.historical_particles))
    logs["learn_from_soup_params"] = {k: v for k, v in soup.params.items() if isinstance(v, (int, float, bool))}

    soup = _load(dill, "results/Soup/soup.dill")
    np.savez_compressed(os.path.join(OUT, "train_soup.npz"), **_pack_particles(soup.historical_particles))
    logs["train_soup_params"] = {k: v for k, v in soup.params.items() if isinstance(v, (int, float, bool))}
    logs["train_soup_counters"] = dict(divergent=0, fix_zero=0, fix_other=13, fix_sec=0, other=7)

    exp = _load(dill, "results/known_fixpoint_variation/experiment.dill")
    np.savez_compressed(os.path.join(OUT, "kfv.npz"), xs=np.asarray(exp.xs, np.float64),
                        ys=np.asarray(exp.ys, np.int64), zs=np.asarray(exp.zs, np.int64))
    logs["kfv"] = dict(trials=exp.trials, max_steps=exp.max_steps, epsilon=exp.epsilon, depth=exp.depth)

    names = ["WeightwiseNeuralNetwork", "AggregatingNeuralNetwork", "RecurrentNeuralNetwork"]
    ac = _load(dill, "results/exp-applying_fixpoint-_1552681870.3570378-0/all_counters.dill")
    logs["applying_fixpoint"] = dict(trials=50, run_count=100, epsilon=1e-4, counters=dict(zip(names, ac)))
    tc = _load(dill, "results/exp-training_fixpoint-_1552658296.0913951-0/all_counters.dill")
    logs["training_fixpoint"] = dict(trials=50, run_count=1000, epsilon=1e-4, counters=dict(zip(names, tc)))
    for key, rel in [("mixed_self_fixpoints", "results/exp-mixed-self-fixpoints-_1552666977.5858653-0/all_data.dill"),
                     ("mixed_soup", "results/exp-mixed-soup-_1552674483.9866457-0/all_data.dill"),
                     ("learn_from_soup", "results/exp-learn-from-soup-_1552658566.5572753-0/all_data.dill")]:
        data = _load(dill, rel)
        nm = _load(dill, rel.replace("all_data", "all_names"))
        logs[key] = [dict(name=str(n), **{k: list(map(float, v)) for k, v in d.items()}) for n, d in zip(nm, data)]

    with open(os.path.join(OUT, "logs.json"), "w") as fh:
        json.dump(logs, fh, indent=1, sort_keys=True)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
