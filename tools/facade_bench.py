"""BASELINE.json config 1 exactly as the reference writes it, through the drop-in CLASS API (one object per net):
100 WeightwiseNeuralNetwork(2,2), 100 x self_attack() each, then the setups' count() -- the loop of
code/setups/applying-fixpoints.py:51-57 -- and the FixpointExperiment.run_net variant (code/network.py:640-646).
Prints one JSON object; bench.py embeds it as its `facade` key."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "self-replicating-neural-networks_b200"))


def count(counters, net):
    if net.is_diverged():
        counters['divergent'] += 1
    elif net.is_fixpoint():
        counters['fix_zero' if net.is_zero() else 'fix_other'] += 1
    elif net.is_fixpoint(2):
        counters['fix_sec'] += 1
    else:
        counters['other'] += 1


def run(nets=100, steps=100):
    import torch
    import network
    import experiment
    network.set_seed(99)
    out = {}
    # warm-up: first launches, arena chunk
    w = network.ParticleDecorator(network.WeightwiseNeuralNetwork(2, 2))
    w.self_attack()
    w.is_fixpoint()
    torch.cuda.synchronize()
    # (a) applying-fixpoints.py: unconditional self_attack loop
    counters = dict(divergent=0, fix_zero=0, fix_other=0, fix_sec=0, other=0)
    t0 = time.perf_counter()
    for _ in range(nets):
        net = network.ParticleDecorator(network.WeightwiseNeuralNetwork(width=2, depth=2).with_keras_params(activation='linear'))
        net.with_params(epsilon=1e-4)
        for _ in range(steps):
            net.self_attack()
        count(counters, net)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    out["self_attack_loop"] = {"nets": nets, "steps": steps, "seconds": dt, "self_applications_per_s": nets * steps / dt,
                               "counters": counters}
    # (b) FixpointExperiment.run_net(net, 100): the engine fuses the loop into one launch per net
    os.chdir(__import__("tempfile").mkdtemp(prefix="srnn_facade_"))
    t0 = time.perf_counter()
    with experiment.FixpointExperiment() as exp:
        for run_id in range(nets):
            net = network.ParticleDecorator(network.WeightwiseNeuralNetwork(width=2, depth=2))
            exp.run_net(net, steps, run_id + 1)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    out["run_net"] = {"nets": nets, "step_limit": steps, "seconds": dt, "nets_per_s": nets / dt, "counters": dict(exp.counters)}
    return out


if __name__ == "__main__":
    print(json.dumps(run()))
