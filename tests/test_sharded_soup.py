"""Sharded PopulationSoup: N ranks == 1 rank, bit for bit (weights, uids, class counts).

CPU (`not gpu`): world_size 2 over gloo, oracle backend -- exercises the product's host-side sharding logic.
GPU: world_size 2 over NCCL with the real engine (skipped on a single-GPU box)."""
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle
from util_parity import assert_same_weights

HERE = os.path.dirname(os.path.abspath(__file__))


def _reference(kind, size, gens):
    """Single-owner oracle run of the same soup (same seeds / params as the worker)."""
    do = oracle.desc(kind, 2, 2, 4 if kind == 1 else 0)
    po = oracle.soup_params(attacking_rate=0.3, learn_from_rate=0.2, train=1, learn_from_severity=2,
                            remove_divergent=True, remove_zero=True, epsilon=1e-4, perm_mode=2)
    w = np.zeros((size, do.n_weights), np.float32)
    oracle.init(do, w, seed=31)
    uid, next_uid = np.arange(size), size
    for g in range(1, gens + 1):
        out = np.zeros_like(w)
        a, _, _ = oracle.soup_sync(do, po, w, size, out, 0, size, 31, g)
        dead = a >= 4
        uid[dead] = next_uid + np.arange(int(dead.sum()))
        next_uid += int(dead.sum())
        w = out
    return w, uid, next_uid, oracle.classify(do, w, 1e-4)[1]


def _run(backend, world, kind, size, gens, tmp_path, exchange="auto", together=0):
    out = str(tmp_path / ("soup_%s_%d_%s_%d.npz" % (backend, kind, exchange, together)))
    port = 29500 + (os.getpid() + kind * 7) % 2000
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(HERE, "dist_soup_worker.py"),
           "--backend", backend, "--out", out, "--kind", str(kind), "--size", str(size), "--gens", str(gens),
           "--exchange", exchange, "--together", str(together)]
    env = dict(os.environ, OMP_NUM_THREADS="2")
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    return np.load(out)


def _check(got, kind, size, gens):
    w, uid, next_uid, counts = _reference(kind, size, gens)
    assert_same_weights(got["w"], w, "sharded vs single owner")
    assert np.array_equal(got["uid"], uid) and int(got["next_uid"]) == next_uid
    assert got["counts"].tolist() == counts.tolist()
    assert next_uid > size                       # some particles were respawned, so uid allocation was exercised


def test_shard_helpers():
    import soup
    assert soup.shard_of(12, 4, 2) == (6, 3)
    with pytest.raises(ValueError):
        soup.shard_of(10, 4, 0)
    assert soup.uid_base_of(0, [3, 5, 2], 100) == (100, 110)
    assert soup.uid_base_of(2, [3, 5, 2], 100) == (108, 110)


@pytest.mark.parametrize("kind", [0, 1])
def test_two_ranks_gloo_equals_one(kind, tmp_path):
    got = _run("gloo", 2, kind, 1024, 3, tmp_path)
    assert int(got["world"]) == 2
    _check(got, kind, 1024, 3)


@pytest.mark.gpu
@pytest.mark.parametrize("kind", [0, 1, 2])
@pytest.mark.parametrize("exchange", ["allgather", "peer", "peer_unchained"])
def test_two_ranks_nccl_equals_one(kind, exchange, tmp_path):
    """all-gather exchange and the fused peer-memory exchange both equal the single-owner result bit for bit."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    got = _run("nccl", 2, kind, 8192, 3, tmp_path, exchange)
    assert int(got["world"]) == 2 and str(got["exchange"]) == exchange
    _check(got, kind, 8192, 3)
    assert int(got["comm_ok"]) == 1      # srnn_comm_allgather (the C ABI's own NCCL communicator) == torch's all-gather


@pytest.mark.gpu
def test_mixed_soup_in_lock_step_nccl(tmp_path):
    """soup.evolve_together: the three kinds of a mixed sharded soup on separate streams with one cross-GPU barrier
    per generation -- every kind still equals its single-owner oracle run bit for bit (weights, uids, counts)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    got = _run("nccl", 2, 0, 8192, 3, tmp_path, "peer", together=1)
    assert int(got["world"]) == 2 and str(got["exchange"]) == "peer"
    for kind in (0, 1, 2):
        sub = {k: got["%s%d" % (k, kind)] for k in ("w", "uid", "counts", "next_uid")}
        assert int(got["time%d" % kind]) == 3
        _check(sub, kind, 8192, 3)


@pytest.mark.gpu
def test_evolve_together_single_gpu_falls_back(tmp_path):
    """Unsharded soups take the plain loop: same result as evolving each soup on its own."""
    import torch
    sys.path.insert(0, os.path.join(os.path.dirname(HERE), "self-replicating-neural-networks_b200"))
    from soup import PopulationSoup, evolve_together
    kw = dict(attacking_rate=0.3, learn_from_rate=0.2, train=1, remove_divergent=True, remove_zero=True, epsilon=1e-4)
    a = [PopulationSoup(k, 4096, aggregates=4 if k == 1 else 0, seed=9, **kw).seed_particles() for k in (0, 1, 2)]
    b = [PopulationSoup(k, 4096, aggregates=4 if k == 1 else 0, seed=9, **kw).seed_particles() for k in (0, 1, 2)]
    evolve_together(a, 3)
    for s in b:
        s.evolve(3)
    for x, y in zip(a, b):
        assert x.time == y.time == 3 and torch.equal(x.w.view(torch.int32), y.w.view(torch.int32))


def test_evolve_together_plain_loop_on_cpu():
    """soup.evolve_together on soups without a peer-memory exchange (here: unsharded, oracle backend) is the plain
    loop: same arenas, uids and generation counters as evolving each soup on its own."""
    import torch
    from dist_soup_worker import OracleBackend
    from soup import PopulationSoup, evolve_together
    kw = dict(attacking_rate=0.3, learn_from_rate=0.2, train=1, remove_divergent=True, remove_zero=True, epsilon=1e-4)

    def make():
        soups = [PopulationSoup(k, 512, aggregates=4 if k == 1 else 0, seed=9, backend=OracleBackend(), **kw) for k in (0, 1)]
        for s in soups:
            s.track_uids = True
            s.seed_particles()
        return soups
    a, b = make(), make()
    evolve_together(a, 3)
    for s in b:
        s.evolve(3)
    for x, y in zip(a, b):
        assert x.time == y.time == 3 and x.next_uid == y.next_uid
        assert torch.equal(x.w.view(torch.int32), y.w.view(torch.int32)) and torch.equal(x.uid, y.uid)
