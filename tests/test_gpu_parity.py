"""GPU parity tests: the CUDA path (through the C ABI of libsrnn_b200.so) against the CPU oracle.

Bar: FP32 weights bit-for-bit (NaN matches NaN), integer outputs (steps, classes, counters, actions) exactly.
Sizes are what the oracle finishes in seconds; the full-size checks use size-independent properties.
"""
import json
import os

import numpy as np
import pytest

import oracle
from util_parity import assert_same_weights, glorot_population

pytestmark = pytest.mark.gpu

SHAPES = [(0, 2, 2, 0), (0, 3, 2, 0), (0, 2, 3, 0), (0, 2, 1, 0), (1, 2, 2, 4), (1, 2, 2, 3), (1, 3, 2, 4),
          (2, 2, 2, 0), (2, 2, 1, 0)]
BENCH_SHAPES = [(0, 2, 2, 0), (1, 2, 2, 4), (2, 2, 2, 0)]


@pytest.fixture(scope="module")
def eng():
    import srnn_engine
    srnn_engine.lib()
    return srnn_engine


def both(eng, key):
    kind, width, depth, aggs = key
    return eng.make_desc(kind, width, depth, aggs), oracle.desc(kind, width, depth, aggs)


def test_descs_agree(eng):
    for key in SHAPES:
        dg, do = both(eng, key)
        assert dg.n_weights == do.n_weights
        assert eng.shape_supported(dg)
    assert eng.shape_supported(eng.make_desc(0, 4, 3)) == 2


# ---------------------------------------------------------------- K1 self-application
@pytest.mark.parametrize("key", SHAPES)
def test_self_apply_fixed_bitwise(eng, key):
    """n_steps unconditional self-applications: weights bit-for-bit, including ragged tile sizes."""
    dg, do = both(eng, key)
    for n, steps in [(1, 1), (255, 3), (257, 2), (5000, 10), (4096, 101)]:
        w = glorot_population(key, n, seed=n + steps)
        pop = eng.Population.from_numpy(dg, w)
        pop.self_apply(steps)
        ref = w.copy()
        oracle.self_apply(do, ref, steps)
        assert_same_weights(pop.numpy(), ref, "%r n=%d steps=%d" % (key, n, steps))


@pytest.mark.parametrize("key", [k for k in SHAPES if k[0] == 0])
def test_self_apply_fixed_zero_signs(eng, key):
    """The fixed-length kernel runs its inner steps in a form that may leave zeros with either sign and only the last
    step in the exact form (csrc/srnn_device.cuh, ww_apply ANYZERO).  Adversarial populations for exactly that: weights
    so small that products and sums flush to zero (of both signs), explicit -0 / +0 entries, a few Inf / NaN rows --
    every step count must equal the oracle's all-exact chain in every bit, signs of zeros included."""
    dg, do = both(eng, key)
    rng = np.random.default_rng(77)
    n, nw = 6000, dg.n_weights
    w = glorot_population(key, n, seed=78)
    scale = np.float32(10.0) ** rng.integers(-24, 1, size=(n, 1)).astype(np.float32)   # 1e-24 .. 1: underflow within 1-3 steps
    w = (w * scale).astype(np.float32)
    w[rng.random((n, nw)) < 0.15] = np.float32(-0.0)
    w[rng.random((n, nw)) < 0.10] = np.float32(0.0)
    w[5, 3], w[6, 0], w[7, nw - 1] = np.inf, np.nan, -np.inf
    saw_negative_zero = False
    for steps in (1, 2, 3, 4, 5, 8, 33):
        pop = eng.Population.from_numpy(dg, w)
        pop.self_apply(steps)
        ref = w.copy()
        oracle.self_apply(do, ref, steps)
        assert_same_weights(pop.numpy(), ref, "%r steps=%d" % (key, steps))
        saw_negative_zero |= bool(((ref == 0) & np.signbit(ref)).any())
    assert saw_negative_zero, "the case is meant to produce legitimate negative zeros as well"


@pytest.mark.parametrize("key", BENCH_SHAPES)
def test_self_apply_strided_records(eng, key):
    """ld > n_weights (padded records) takes the un-staged path."""
    dg, do = both(eng, key)
    n = 1000
    w = glorot_population(key, n, seed=4)
    pad = np.zeros((n, dg.n_weights + 3), np.float32)
    pad[:, :dg.n_weights] = w
    pop = eng.Population.from_numpy(dg, pad)
    pop.self_apply(7)
    ref = w.copy()
    oracle.self_apply(do, ref, 7)
    assert_same_weights(pop.numpy(), ref)
    assert (pop.w[:, dg.n_weights:].cpu().numpy() == 0).all()


@pytest.mark.parametrize("key", SHAPES)
@pytest.mark.parametrize("eps", [1e-14, 1e-4])
def test_run_net_until_and_classes(eng, key, eps):
    """run_net stop rule + fused classification: steps, classes, counters exact; weights and trajectories bitwise."""
    dg, do = both(eng, key)
    n = 3000
    w = glorot_population(key, n, seed=7)
    w[0] = 0.0
    pop = eng.Population.from_numpy(dg, w)
    out = pop.self_apply(100, mode=eng.RUN_UNTIL, eps=eps, want_steps=True, want_cls=True, want_counters=True,
                         want_traj=(key in BENCH_SHAPES))
    ref = w.copy()
    if key in BENCH_SHAPES:
        steps, traj = oracle.self_apply(do, ref, 100, mode=oracle.RUN_UNTIL, eps=eps, traj=True)
    else:
        steps = oracle.self_apply(do, ref, 100, mode=oracle.RUN_UNTIL, eps=eps)
    cls, cnt = oracle.classify(do, ref, eps)
    assert np.array_equal(out["steps"].cpu().numpy(), steps)
    assert_same_weights(pop.numpy(), ref)
    assert np.array_equal(out["cls"].cpu().numpy(), cls)
    assert out["counters"].cpu().numpy().tolist() == cnt.tolist()
    if key in BENCH_SHAPES:
        gt = out["traj"].cpu().numpy()
        for i in range(0, n, 37):
            assert_same_weights(gt[:steps[i], i, :dg.n_weights], traj[:steps[i], i], "traj %d" % i)


def test_golden_ww_run_net(eng, golden_dir):
    """The reference's own 20 WW(2,2) trajectories, free-running on the GPU from the golden initial weights:
    counters {11 divergent, 9 fix_zero}, saved-state counts and every stored state bit-for-bit."""
    g = np.load(os.path.join(golden_dir, "ww_self_application.npz"))
    W, P = g["weights"], g["particle"]
    uids = np.unique(P)
    init = np.ascontiguousarray(np.stack([W[P == u][0] for u in uids]))
    dg, _ = both(eng, (0, 2, 2, 0))
    pop = eng.Population.from_numpy(dg, init)
    out = pop.self_apply(100, mode=eng.RUN_UNTIL, eps=1e-14, want_steps=True, want_counters=True, want_traj=True)
    assert out["counters"].cpu().numpy().tolist() == [11, 9, 0, 0, 0]
    steps, traj = out["steps"].cpu().numpy(), out["traj"].cpu().numpy()
    for i, u in enumerate(uids):
        S = W[P == u]
        saved = 1 + sum(np.isfinite(traj[s, i]).all() for s in range(steps[i]))
        assert saved == len(S)
        for s in range(1, len(S)):
            assert_same_weights(traj[s - 1, i], S[s], "golden particle %d step %d" % (u, s))


def test_golden_agg_teacher_forced(eng, golden_dir):
    g = np.load(os.path.join(golden_dir, "agg_self_application.npz"))
    W, P, T = g["weights"], g["particle"], g["time"]
    pairs = [(i, i + 1) for i in range(len(P) - 1) if P[i] == P[i + 1] and T[i + 1] == T[i] + 1]
    dg, _ = both(eng, (1, 2, 2, 4))
    pop = eng.Population.from_numpy(dg, W[[p[0] for p in pairs]])
    pop.self_apply(1)
    assert_same_weights(pop.numpy(), W[[p[1] for p in pairs]])


# ---------------------------------------------------------------- K2/K3/K4 attack pairs
@pytest.mark.parametrize("key", SHAPES)
def test_apply_pairs(eng, key):
    dg, do = both(eng, key)
    n = 2048
    w = glorot_population(key, n, seed=9)
    rng = np.random.default_rng(1)
    victims = rng.permutation(n)[: n // 2]
    rest = np.setdiff1d(np.arange(n), victims)
    attackers = rng.choice(rest, size=victims.size)
    pop = eng.Population.from_numpy(dg, w)
    pop.attack(attackers, victims)
    ref = w.copy()
    oracle.apply(do, ref, attackers, ref, victims, ref, victims, victims.size)
    assert_same_weights(pop.numpy(), ref)
    # self-attack through the pair path (op = in = out)
    pop.attack(rest[:100], rest[:100])
    oracle.apply(do, ref, rest[:100], ref, rest[:100], ref, rest[:100], 100)
    assert_same_weights(pop.numpy(), ref)


# ---------------------------------------------------------------- K6 classify / predicates
@pytest.mark.parametrize("key", SHAPES)
def test_classify_and_flags(eng, key):
    dg, do = both(eng, key)
    n = 4000
    w = glorot_population(key, n, seed=12)
    oracle.self_apply(do, w, 3)            # a mix of shrinking, growing and non-finite nets
    w[1] = 0.0
    w[2, 0] = np.inf
    w[3, -1] = np.nan
    if key == (0, 2, 2, 0):
        w[4] = 0.0
        w[4, [0, 8, 12]] = 1.0              # the identity fixpoint: fix_other
    for eps in (1e-14, 1e-4, 0.5):
        pop = eng.Population.from_numpy(dg, w)
        for fs in (True, False):
            cls, cnt, flags = pop.classify(eps, with_fix_sec=fs, want_flags=True)
            rcls, rcnt = oracle.classify(do, w, eps, with_fix_sec=fs)
            assert np.array_equal(cls.cpu().numpy(), rcls)
            assert cnt.cpu().numpy().tolist() == rcnt.tolist()
        assert np.array_equal(flags.cpu().numpy(), oracle.predicates(do, w, eps))


# ---------------------------------------------------------------- K7 init
@pytest.mark.parametrize("key", SHAPES)
def test_init_bitwise(eng, key):
    dg, do = both(eng, key)
    n = 3000
    pop = eng.Population(dg, n)
    pop.init(seed=1234, slot0=77, generation=5)
    ref = np.zeros((n, do.n_weights), np.float32)
    oracle.init(do, ref, seed=1234, slot0=77, generation=5)
    assert_same_weights(pop.numpy(), ref)
    idx = np.array([5, 17, 2999])
    pop.init(seed=99, slot0=0, generation=1, idx=idx)
    oracle.init(do, ref, seed=99, slot0=0, generation=1, idx=idx)
    assert_same_weights(pop.numpy(), ref)


# ---------------------------------------------------------------- K5 train / learn_from
@pytest.mark.parametrize("key", SHAPES)
@pytest.mark.parametrize("perm_mode", [0, 1, 2])
def test_train_bitwise(eng, key, perm_mode):
    dg, do = both(eng, key)
    n, epochs = 700, 12
    w = glorot_population(key, n, seed=21)
    ns = dg.n_weights if key[0] == 0 else 1
    perm = None
    if perm_mode == 1:
        rs = np.random.RandomState(5)
        perm = np.stack([rs.permutation(ns) for _ in range(epochs)]).astype(np.uint8)
    pop = eng.Population.from_numpy(dg, w)
    loss = pop.train(epochs, lr=0.01, perm_mode=perm_mode, perm=perm, seed=3, slot0=10, epoch0=4)
    ref = w.copy()
    rloss = oracle.train(do, ref, epochs, lr=0.01, perm_mode=perm_mode, perm=perm, seed=3, slot0=10, epoch0=4)
    assert_same_weights(pop.numpy(), ref, "train %r perm=%d" % (key, perm_mode))
    gl = loss.cpu().numpy()
    assert np.array_equal(np.isnan(gl), np.isnan(rloss))
    m = ~np.isnan(gl)
    assert np.array_equal(gl[m], rloss[m])


@pytest.mark.parametrize("key", SHAPES)
def test_learn_from_bitwise(eng, key):
    dg, do = both(eng, key)
    n = 600
    w = glorot_population(key, n, seed=31)
    t = glorot_population(key, n, seed=32)
    tidx = np.random.default_rng(2).integers(0, n, n)
    pop, tea = eng.Population.from_numpy(dg, w), eng.Population.from_numpy(dg, t)
    pop.train(5, perm_mode=2, seed=8, teacher=tea, teacher_idx=tidx)
    ref = w.copy()
    oracle.train(do, ref, 5, perm_mode=2, seed=8, teacher_w=t, teacher_idx=tidx)
    assert_same_weights(pop.numpy(), ref)


def test_training_reaches_fixpoints(eng):
    """training-fixpoints.py statistic at population scale: 1000 epochs -> WW all fix_other, AGG all other (eps 1e-4)."""
    dg, _ = both(eng, (0, 2, 2, 0))
    pop = eng.Population(dg, 10000).init(seed=5)
    pop.train(1000, perm_mode=2, seed=6, want_loss=False)
    cnt = pop.classify(1e-4)[1].cpu().numpy()
    assert cnt[2] >= 9700, cnt
    dg, _ = both(eng, (1, 2, 2, 4))
    pop = eng.Population(dg, 10000).init(seed=5)
    pop.train(1000, want_loss=False)
    cnt = pop.classify(1e-4)[1].cpu().numpy()
    assert cnt[4] >= 9000, cnt


# ---------------------------------------------------------------- K9 known-fixpoint variation
def test_kfv_bitwise_and_statistics(eng, golden_dir):
    dg, do = both(eng, (0, 2, 2, 0))
    base = np.zeros(14, np.float32)
    base[[0, 8, 12]] = 1.0
    g = np.load(os.path.join(golden_dir, "kfv.npz"))
    scale = 1.0
    for level in range(10):
        n = 20000
        out = eng.kfv(dg, base, n, scale, level, 100, 1e-4, seed=99, slot0=5, want_weights=True)
        ys, zs, wr = oracle.kfv(do, base, n, scale, level, 100, 1e-4, seed=99, slot0=5, want_weights=True)
        assert np.array_equal(out["ys"].cpu().numpy(), ys)
        assert np.array_equal(out["zs"].cpu().numpy(), zs)
        assert_same_weights(out["w"].cpu().numpy(), wr)
        hist = out["hist"].cpu().numpy()
        assert np.array_equal(hist[0], np.bincount(ys, minlength=101))
        assert np.array_equal(hist[1], np.bincount(zs, minlength=101))
        # without the weights the kernel may leave zeros of either sign inside its loop: the counts must not notice
        lean = eng.kfv(dg, base, n, scale, level, 100, 1e-4, seed=99, slot0=5)
        assert lean["w"] is None
        assert np.array_equal(lean["ys"].cpu().numpy(), ys) and np.array_equal(lean["zs"].cpu().numpy(), zs)
        assert np.array_equal(lean["hist"].cpu().numpy(), hist)
        ry = g["ys"][level * 100:(level + 1) * 100]
        assert abs(ys.mean() - ry.mean()) < 4 * max(ry.std(), 0.3) / 10 + 0.1
        scale /= 10.0


# ---------------------------------------------------------------- K8 soup
@pytest.mark.parametrize("key", BENCH_SHAPES)
def test_soup_ordered_bitwise(eng, key):
    """Exact-order Soup.evolve: a few generations with attacks, learn_from, training and culling."""
    dg, do = both(eng, key)
    n = 40
    w = glorot_population(key, n, seed=41)
    w[3] = 0.0
    rng = np.random.default_rng(6)
    pg = eng.make_soup_params(attacking_rate=0.3, learn_from_rate=0.3, train=2, learn_from_severity=2,
                              remove_divergent=True, remove_zero=True, epsilon=1e-4, perm_mode=2)
    po = oracle.soup_params(attacking_rate=0.3, learn_from_rate=0.3, train=2, learn_from_severity=2,
                            remove_divergent=True, remove_zero=True, epsilon=1e-4, perm_mode=2)
    pop = eng.Population.from_numpy(dg, w)
    ref = w.copy()
    for gen in range(1, 6):
        att = np.where(rng.random(n) < 0.3, rng.integers(0, n, n), -1)
        lrn = np.where(rng.random(n) < 0.3, rng.integers(0, n, n), -1)
        action, loss, logged = pop.soup_ordered(pg, att, lrn, seed=17, gen=gen, want_logged=True)
        ra, rl, rlog = oracle.soup_ordered(do, po, ref, att, lrn, seed=17, gen=gen, want_logged=True)
        assert_same_weights(pop.numpy(), ref, "gen %d" % gen)
        assert np.array_equal(action.cpu().numpy(), ra)
        assert_same_weights(logged.cpu().numpy(), rlog)
        gl = loss.cpu().numpy()
        assert np.array_equal(np.isnan(gl), np.isnan(rl)) and np.array_equal(gl[~np.isnan(gl)], rl[~np.isnan(rl)])


@pytest.mark.parametrize("key", BENCH_SHAPES)
@pytest.mark.parametrize("train,sev", [(0, 1), (2, 3)])
@pytest.mark.parametrize("path", [1, 2])
def test_soup_sync_bitwise_and_shard_invariance(eng, key, train, sev, path):
    """Snapshot-synchronous generations: one owner of all particles == the oracle, and two owners of half the
    population each (the multi-GPU sharding, emulated on one device) produce the same arena."""
    import torch
    dg, do = both(eng, key)
    n = 3000
    kw = dict(attacking_rate=0.4, learn_from_rate=0.2, train=train, learn_from_severity=sev, remove_divergent=True,
              remove_zero=True, epsilon=1e-4, perm_mode=2)
    pg, po = eng.make_soup_params(engine_path=path, **kw), oracle.soup_params(**kw)   # fused kernel / multi-pass
    w = glorot_population(key, n, seed=51)
    snap = torch.from_numpy(w).cuda()
    ref = w.copy()
    scratch = torch.zeros(eng.soup_scratch_len(n, n), dtype=torch.int32, device="cuda")
    for gen in range(1, 4):
        out = torch.zeros_like(snap)
        action = torch.zeros(n, dtype=torch.uint8, device="cuda")
        dead = torch.zeros(n, dtype=torch.uint8, device="cuda")
        eng.soup_sync(dg, pg, snap, n, out, 0, n, 23, gen, scratch, action=action, dead=dead)
        rout = np.zeros_like(ref)
        ra, rl, natt = oracle.soup_sync(do, po, ref, n, rout, 0, n, 23, gen)
        assert natt > 0.3 * n
        assert_same_weights(out.cpu().numpy(), rout, "gen %d" % gen)
        assert np.array_equal(action.cpu().numpy(), ra)
        # sharded: two owners
        half = n // 2 + 7
        o2 = torch.zeros_like(snap)
        eng.soup_sync(dg, pg, snap, n, o2[:half], 0, half, 23, gen, scratch)
        eng.soup_sync(dg, pg, snap, n, o2[half:], half, n - half, 23, gen, scratch)
        assert torch.equal(o2.view(torch.int32), out.view(torch.int32))
        # peer-memory variant: the two halves live in separate arenas and partner records are fetched by pointer
        sh = [snap[:n // 2].clone(), snap[n // 2:].clone()]
        o3 = torch.zeros_like(snap)
        for r in range(2):
            eng.soup_sync_peers(dg, pg, [t.data_ptr() for t in sh], n // 2, r, sh[0].shape[1],
                                o3[r * (n // 2):(r + 1) * (n // 2)], 23, gen, scratch)
        assert torch.equal(o3.view(torch.int32), out.view(torch.int32))
        # uid allocation for the respawned particles
        uid = torch.arange(n, dtype=torch.int64, device="cuda")
        tmp = torch.zeros((n + 1023) // 1024 + 1, dtype=torch.int64, device="cuda")
        nd = torch.zeros(1, dtype=torch.int64, device="cuda")
        eng.assign_uids(dead, uid, 10 ** 6, tmp, nd)
        dn = dead.cpu().numpy().astype(bool)
        want = np.arange(n)
        want[dn] = 10 ** 6 + np.arange(dn.sum())
        assert np.array_equal(uid.cpu().numpy(), want) and int(nd.item()) == int(dn.sum())
        snap, ref = out, rout


@pytest.mark.parametrize("key", BENCH_SHAPES)
@pytest.mark.parametrize("world", [2, 3])
def test_soup_chain_peers_on_one_gpu(eng, key, world):
    """srnn_soup_chain_peers / srnn_soup_push_lists, the sharded generation whose attack lists are built by the
    attackers' owners: `world` shards emulated on one GPU (separate arenas and list arrays, every pointer a "peer"),
    4 generations == the single-owner oracle run, bit for bit; the list heads are left clean for the next run."""
    import torch
    dg, do = both(eng, key)
    shard = 1100
    n = shard * world
    kw = dict(attacking_rate=0.4, learn_from_rate=0.2, train=0, learn_from_severity=1, remove_divergent=True,
              remove_zero=True, epsilon=1e-4, perm_mode=2)
    pg, po = eng.make_soup_params(**kw), oracle.soup_params(**kw)
    w = glorot_population(key, n, seed=77)
    ref = w.copy()
    cur = [torch.from_numpy(w[r * shard:(r + 1) * shard].copy()).cuda() for r in range(world)]
    nxt = [torch.zeros_like(t) for t in cur]
    lists = [torch.full((2, shard + n), -1, dtype=torch.int32, device="cuda") for _ in range(world)]

    def tables(par):
        heads = eng.shard_ptr_array([t[par].data_ptr() for t in lists])
        nexts = eng.shard_ptr_array([t[par].data_ptr() + 4 * shard for t in lists])
        return heads, nexts
    h1, n1 = tables(1)
    for r in range(world):
        eng.soup_push_lists(pg, world, shard, r, 23, 1, h1, n1)
    act = [torch.zeros(shard, dtype=torch.uint8, device="cuda") for _ in range(world)]
    flag_mem = [torch.zeros(world + 1, dtype=torch.int32, device="cuda") for _ in range(world)]
    flags = eng.shard_ptr_array([t.data_ptr() for t in flag_mem])
    for gen in range(1, 5):
        (hc, nc), (hn, nn) = tables(gen & 1), tables((gen + 1) & 1)
        for r in range(world):
            # the fused cross-GPU signal: wait (on the device) for generation gen-1 of every rank, signal gen when done
            eng.soup_chain_peers(dg, pg, [t.data_ptr() for t in cur], shard, r, cur[0].shape[1], nxt[r], 23, gen, hc, nc,
                                 hn if gen < 4 else None, nn if gen < 4 else None, action=act[r], flags=flags,
                                 wait_gen=gen - 1)
        for t in flag_mem:
            assert t[:world].tolist() == [gen] * world
        rout = np.zeros_like(ref)
        ra, _, natt = oracle.soup_sync(do, po, ref, n, rout, 0, n, 23, gen)
        assert natt > 0.3 * n
        got = torch.cat(nxt).cpu().numpy()
        assert_same_weights(got, rout, "gen %d" % gen)
        assert np.array_equal(torch.cat(act).cpu().numpy(), ra)
        cur, nxt, ref = nxt, cur, rout
    for t in lists:
        assert int((t[:, :shard] != -1).sum().item()) == 0       # every consumed head was reset


@pytest.mark.parametrize("key", [(0, 2, 2, 0), (2, 2, 2, 0)])
def test_soup_deferred_learners(eng, key):
    """Large weightwise and recurrent soups hand their learn_from epochs to a dense second kernel
    (soup_deferred_learners_kernel; taken from 2^20 particles per launch on, generations without self-training).  Same
    results as the oracle's generation, bit for bit -- weights, actions, respawns -- through srnn_soup_sync (one owner
    and two uneven owners) and through the chained sharded entry srnn_soup_chain_peers."""
    import torch
    dg, do = both(eng, key)
    shard = (1 << 20) + 64
    n = 2 * shard
    kw = dict(attacking_rate=0.2, learn_from_rate=0.2, train=0, learn_from_severity=2, remove_divergent=True,
              remove_zero=True, epsilon=1e-4, perm_mode=2)
    pg, po = eng.make_soup_params(**kw), oracle.soup_params(**kw)
    w = glorot_population(key, n, seed=91)
    ref = w.copy()
    snap = torch.from_numpy(w).cuda()
    scratch = torch.zeros(eng.soup_scratch_len(n, n), dtype=torch.int32, device="cuda")
    lists = [torch.full((2, shard + n), -1, dtype=torch.int32, device="cuda") for _ in range(2)]
    small = [torch.zeros(shard + 16, dtype=torch.int32, device="cuda") for _ in range(2)]

    def tables(par):
        return (eng.shard_ptr_array([t[par].data_ptr() for t in lists]),
                eng.shard_ptr_array([t[par].data_ptr() + 4 * shard for t in lists]))
    h1, n1 = tables(1)
    for r in range(2):
        eng.soup_push_lists(pg, 2, shard, r, 23, 1, h1, n1)
    for gen in (1, 2):
        out = torch.zeros_like(snap)
        action = torch.zeros(n, dtype=torch.uint8, device="cuda")
        dead = torch.zeros(n, dtype=torch.uint8, device="cuda")
        eng.soup_sync(dg, pg, snap, n, out, 0, n, 23, gen, scratch, action=action, dead=dead)
        rout = np.zeros_like(ref)
        ra, _, _ = oracle.soup_sync(do, po, ref, n, rout, 0, n, 23, gen)
        assert_same_weights(out.cpu().numpy(), rout, "gen %d" % gen)
        assert np.array_equal(action.cpu().numpy(), ra)
        assert np.array_equal(dead.cpu().numpy(), (ra >= 4).astype(np.uint8)) and (ra == 2).sum() > 0.1 * n
        # two uneven owners (the second one is below the threshold: it runs its learners inside the tile kernel)
        cut = n - 5000
        o2 = torch.zeros_like(snap)
        eng.soup_sync(dg, pg, snap, n, o2[:cut], 0, cut, 23, gen, scratch)
        eng.soup_sync(dg, pg, snap, n, o2[cut:], cut, n - cut, 23, gen, scratch)
        assert torch.equal(o2.view(torch.int32), out.view(torch.int32))
        # chained sharded entry, two shards emulated on this GPU
        cur = [snap[:shard].clone(), snap[shard:].clone()]
        o3 = torch.zeros_like(snap)
        (hc, nc), (hn, nn) = tables(gen & 1), tables((gen + 1) & 1)
        for r in range(2):
            eng.soup_chain_peers(dg, pg, [t.data_ptr() for t in cur], shard, r, cur[0].shape[1],
                                 o3[r * shard:(r + 1) * shard], 23, gen, hc, nc, hn, nn, scratch=small[r])
        assert torch.equal(o3.view(torch.int32), out.view(torch.int32))
        snap, ref = out, rout


@pytest.mark.parametrize("key", BENCH_SHAPES)
@pytest.mark.parametrize("perm_mode", [2, 0])
def test_soup_run_small_training_soup(eng, key, perm_mode):
    """Small training soups (one wave of CTAs) run one particle per thread, weightwise ones with the self-training kernel's
    unrolled epoch: several epochs per generation (loss-less epochs + the reported last one), Philox and fixed sample
    order, through the chained run -- against the oracle's generations."""
    import torch
    dg, do = both(eng, key)
    n = 3000
    kw = dict(attacking_rate=0.3, learn_from_rate=0.2, train=4, learn_from_severity=2, remove_divergent=True,
              remove_zero=True, epsilon=1e-4, perm_mode=perm_mode)
    pg, po = eng.make_soup_params(**kw), oracle.soup_params(**kw)
    w = glorot_population(key, n, seed=61)
    ref = w.copy()
    a = torch.from_numpy(w).cuda()
    b = torch.zeros_like(a)
    scratch = torch.zeros(eng.soup_scratch_len(n, n), dtype=torch.int32, device="cuda")
    gens = 3
    for gen in range(1, gens + 1):
        rout = np.zeros_like(ref)
        oracle.soup_sync(do, po, ref, n, rout, 0, n, 17, gen)
        ref = rout
    eng.soup_sync_run(dg, pg, a, b, 17, 1, gens, scratch)
    assert_same_weights((b if gens % 2 else a).cpu().numpy(), ref, "%r perm_mode %d" % (key, perm_mode))
    # one generation through the per-call entry, with the loss of the last epoch
    out = torch.zeros_like(a)
    loss = torch.zeros(n, dtype=torch.float64, device="cuda")
    start = torch.from_numpy(w).cuda()
    eng.soup_sync(dg, pg, start, n, out, 0, n, 17, 1, scratch, loss=loss)
    rout = np.zeros_like(w)
    _, rl, _ = oracle.soup_sync(do, po, w, n, rout, 0, n, 17, 1)
    assert_same_weights(out.cpu().numpy(), rout)
    gl = loss.cpu().numpy()
    assert np.array_equal(np.isnan(gl), np.isnan(rl)) and np.array_equal(gl[~np.isnan(gl)], rl[~np.isnan(rl)])


@pytest.mark.parametrize("key", SHAPES)
@pytest.mark.parametrize("n", [1, 255, 256, 257, 1000])
def test_soup_sync_every_fast_shape_and_ragged_tiles(eng, key, n):
    """Every compiled shape (the tile holds 128 or 256 particles depending on the record size) and particle counts
    on both sides of the tile boundaries, through both the per-generation call and the chained run."""
    import torch
    dg, do = both(eng, key)
    kw = dict(attacking_rate=0.5, learn_from_rate=0.3, train=1, learn_from_severity=1, remove_divergent=True,
              remove_zero=True, epsilon=1e-4, perm_mode=2)
    pg, po = eng.make_soup_params(**kw), oracle.soup_params(**kw)
    w = glorot_population(key, n, seed=5)
    ref = w.copy()
    snap = torch.from_numpy(w).cuda()
    a, b = snap.clone(), torch.zeros_like(snap)
    scratch = torch.zeros(eng.soup_scratch_len(n, n), dtype=torch.int32, device="cuda")
    gens = 3
    for gen in range(1, gens + 1):
        out = torch.zeros_like(snap)
        eng.soup_sync(dg, pg, snap, n, out, 0, n, 9, gen, scratch)
        rout = np.zeros_like(ref)
        oracle.soup_sync(do, po, ref, n, rout, 0, n, 9, gen)
        assert_same_weights(out.cpu().numpy(), rout, "n %d gen %d" % (n, gen))
        snap, ref = out, rout
    scratch.zero_()
    eng.soup_sync_run(dg, pg, a, b, 9, 1, gens, scratch)
    last = b if gens % 2 else a
    assert torch.equal(last.view(torch.int32), snap.view(torch.int32))


# ---------------------------------------------------------------- full-size, size-independent properties
def test_full_size_properties(eng):
    """2^22 WW(2,2) nets x 100 steps (the bench workload's shape): (i) the identity fixpoint is invariant,
    (ii) two runs are bit-identical, (iii) a strided 1/1024 sample equals the oracle, (iv) class counts of the
    whole population equal the sum over shards."""
    import torch
    dg, do = both(eng, (0, 2, 2, 0))
    n = 1 << 22
    pop = eng.Population(dg, n).init(seed=0)
    pop.w[::4096] = 0.0
    pop.w[::4096, 0] = 1.0
    pop.w[::4096, 8] = 1.0
    pop.w[::4096, 12] = 1.0
    start = pop.w.clone()
    pop.self_apply(100)
    first = pop.w.clone()
    assert torch.equal(first[::4096].view(torch.int32), start[::4096].view(torch.int32))
    pop.w.copy_(start)
    pop.self_apply(100)
    assert torch.equal(pop.w.view(torch.int32), first.view(torch.int32))
    sample = np.ascontiguousarray(start[5::1024].cpu().numpy())
    oracle.self_apply(do, sample, 100)
    assert_same_weights(first[5::1024].cpu().numpy(), sample)
    cls, cnt = pop.classify(1e-4)
    parts = sum(eng.Population(dg, n // 4, w=first[q * (n // 4):(q + 1) * (n // 4)]).classify(1e-4)[1] for q in range(4))
    assert torch.equal(cnt, parts) and int(cnt.sum()) == n
    rc, rcnt = oracle.classify(do, sample, 1e-4)
    assert np.array_equal(cls[5::1024].cpu().numpy(), rc)


# ---------------------------------------------------------------- edge cases of the C ABI
def test_empty_and_invalid_calls(eng):
    """n = 0 is a no-op; bad arguments are status codes with a message, never a crash or a silent fallback."""
    import ctypes as C
    import torch
    lib = eng.lib()
    dg, _ = both(eng, (0, 2, 2, 0))
    w = torch.zeros((4, 14), dtype=torch.float32, device="cuda")
    p = C.c_void_p(w.data_ptr())
    assert lib.srnn_self_apply(C.byref(dg), p, C.c_int64(14), C.c_int64(0), C.c_int32(5), C.c_int32(0), C.c_double(1e-4),
                               C.c_int32(1), None, None, None, None, None) == 0
    assert lib.srnn_classify(C.byref(dg), p, C.c_int64(14), C.c_int64(0), C.c_double(1e-4), C.c_int32(1), None, None, None,
                             None) == 0
    # record stride smaller than the record
    assert lib.srnn_self_apply(C.byref(dg), p, C.c_int64(13), C.c_int64(4), C.c_int32(1), C.c_int32(0), C.c_double(1e-4),
                               C.c_int32(1), None, None, None, None, None) == 1
    assert b"stride" in lib.srnn_last_error_string()
    # NULL arena
    assert lib.srnn_self_apply(C.byref(dg), None, C.c_int64(14), C.c_int64(4), C.c_int32(1), C.c_int32(0), C.c_double(1e-4),
                               C.c_int32(1), None, None, None, None, None) == 1
    # epsilon below FLT_MIN is refused where predicates are evaluated
    cls = torch.zeros(4, dtype=torch.uint8, device="cuda")
    assert lib.srnn_classify(C.byref(dg), p, C.c_int64(14), C.c_int64(4), C.c_double(1e-60), C.c_int32(1),
                             C.c_void_p(cls.data_ptr()), None, None, None) == 1
    assert b"epsilon" in lib.srnn_last_error_string()
    # unknown run mode / negative step count
    assert lib.srnn_self_apply(C.byref(dg), p, C.c_int64(14), C.c_int64(4), C.c_int32(1), C.c_int32(7), C.c_double(1e-4),
                               C.c_int32(1), None, None, None, None, None) == 1
    assert lib.srnn_self_apply(C.byref(dg), p, C.c_int64(14), C.c_int64(4), C.c_int32(-1), C.c_int32(0), C.c_double(1e-4),
                               C.c_int32(1), None, None, None, None, None) == 1
    # a shape outside SRNN_SHAPES: per-net operations, snapshot soups and the KFV sweep work (generic path); the
    # exact-order and single-thread soup entries exist for the instantiated shapes only and say so
    d2 = eng.make_desc(0, 4, 3)
    w2 = torch.zeros((4, d2.n_weights), dtype=torch.float32, device="cuda")
    assert lib.srnn_self_apply(C.byref(d2), C.c_void_p(w2.data_ptr()), C.c_int64(d2.n_weights), C.c_int64(4), C.c_int32(1),
                               C.c_int32(0), C.c_double(1e-4), C.c_int32(1), None, None, None, None, None) == 0
    base = np.zeros(d2.n_weights, np.float32)
    assert lib.srnn_kfv(C.byref(d2), base.ctypes.data_as(C.c_void_p), C.c_int64(4), C.c_double(0.1), C.c_uint32(0),
                        C.c_int32(10), C.c_double(1e-4), C.c_uint64(0), C.c_uint64(0), None, None, None, None,
                        C.c_int64(d2.n_weights), None) == 0
    sp = eng.make_soup_params()
    att = torch.full((4,), -1, dtype=torch.int64, device="cuda")
    assert lib.srnn_soup_ordered(C.byref(d2), C.byref(sp), C.c_void_p(w2.data_ptr()), C.c_int64(d2.n_weights), C.c_int64(4),
                                 C.c_void_p(att.data_ptr()), C.c_void_p(att.data_ptr()), C.c_uint64(0), C.c_uint64(1), None,
                                 None, None, None, C.c_int64(0), C.c_int32(0), None, None) == 2
    assert b"SRNN_SHAPES" in lib.srnn_last_error_string()
    torch.cuda.synchronize()
    assert float(w.abs().sum()) == 0.0


def test_nan_inf_are_states_not_errors(eng):
    """A NaN attacker infects its victim; infinities propagate with IEEE semantics."""
    dg, do = both(eng, (0, 2, 2, 0))
    w = glorot_population((0, 2, 2, 0), 64, seed=3)
    w[0, 3] = np.nan
    w[1, 0] = np.inf
    w[2, 13] = -np.inf
    pop = eng.Population.from_numpy(dg, w)
    pop.attack(np.array([0, 1, 2]), np.array([10, 11, 12]))
    ref = w.copy()
    oracle.apply(do, ref, [0, 1, 2], ref, [10, 11, 12], ref, [10, 11, 12], 3)
    assert_same_weights(pop.numpy(), ref)
    assert not np.isfinite(pop.numpy()[10:13]).all(axis=1).any()
    cls = pop.classify(1e-4)[0].cpu().numpy()
    assert (cls[[0, 1, 2, 10, 11, 12]] == 0).all()


@pytest.mark.parametrize("key", [(1, 2, 2, 4), (1, 2, 2, 3), (1, 3, 2, 4)])
def test_agg_block_means_exact_division(eng, key):
    """The AGG block mean divides a float64 sum by a compile-time count with a three-instruction correctly rounded
    quotient (div_count); the oracle uses the IEEE division.  Victims with wildly mixed exponents, cancellation, huge
    and tiny magnitudes, infinities and NaNs must come out bit-identical (counts 5, 6+, 3: non powers of two)."""
    dg, do = both(eng, key)
    nw = dg.n_weights
    n = 1 << 16
    rng = np.random.default_rng(17)
    bits = rng.integers(0, 2 ** 32, (n, nw), dtype=np.uint64).astype(np.uint32)
    w = bits.view(np.float32).copy()                      # every exponent, both signs, NaNs, infinities
    w[np.abs(w) < np.float32(1.1754944e-38)] = 0.0        # no denormals: stored weights are always flushed results (R1)
    w[: n // 4] = (rng.standard_normal((n // 4, nw)) * 10.0 ** rng.integers(-30, 30, (n // 4, 1))).astype(np.float32)
    w[n // 4: n // 2] = glorot_population(key, n // 4, seed=3)
    w[n // 2, :] = np.float32(3.4028235e38)              # the sum leaves the float32 range, the mean returns to it
    w[n // 2 + 1, :] = np.float32(1.1754944e-38)
    w[n // 2 + 2, :] = 0.0
    w[n // 2 + 3, :] = -0.0
    ops = glorot_population(key, n, seed=4)               # finite operators: the result shows the means, not just NaN
    pop_op, pop_in, pop_out = (eng.Population.from_numpy(dg, a) for a in (ops, w, np.zeros_like(w)))
    pop_op.apply_pairs(None, pop_in, None, pop_out, None, n=n)
    ref = np.zeros_like(w)
    ar = np.arange(n)
    oracle.apply(do, ops, ar, w, ar, ref, ar, n)
    assert_same_weights(pop_out.numpy(), ref, "agg means")
    assert np.isfinite(ref).all(axis=1).sum() > n // 2


# ---------------------------------------------------------------- generic run-time-shape path
GENERIC_SHAPES = [(0, 4, 3, 0), (0, 1, 1, 0), (0, 5, 2, 0), (1, 3, 3, 5), (1, 4, 1, 2), (2, 3, 2, 0), (2, 1, 3, 0)]


@pytest.mark.parametrize("key", GENERIC_SHAPES)
def test_generic_shapes_bitwise(eng, key):
    """Shapes outside SRNN_SHAPES take the run-time-shape kernels: init, self_apply (fixed / run_net / trajectory /
    classes), classify + flags, attack pairs, forward, train and learn_from -- all bit-for-bit equal to the oracle."""
    dg, do = both(eng, key)
    assert eng.shape_supported(dg) == 2
    n = 600
    pop = eng.Population(dg, n).init(seed=8, slot0=3, generation=2)
    ref = np.zeros((n, do.n_weights), np.float32)
    oracle.init(do, ref, seed=8, slot0=3, generation=2)
    assert_same_weights(pop.numpy(), ref, "init")
    w = glorot_population(key, n, seed=14)
    pop = eng.Population.from_numpy(dg, w)
    pop.self_apply(6)
    ref = w.copy()
    oracle.self_apply(do, ref, 6)
    assert_same_weights(pop.numpy(), ref, "self_apply fixed")
    pop = eng.Population.from_numpy(dg, w)
    out = pop.self_apply(50, mode=eng.RUN_UNTIL, eps=1e-4, want_steps=True, want_cls=True, want_counters=True, want_traj=True)
    ref = w.copy()
    steps, traj = oracle.self_apply(do, ref, 50, mode=oracle.RUN_UNTIL, eps=1e-4, traj=True)
    cls, cnt = oracle.classify(do, ref, 1e-4)
    assert np.array_equal(out["steps"].cpu().numpy(), steps) and np.array_equal(out["cls"].cpu().numpy(), cls)
    assert out["counters"].cpu().numpy().tolist() == cnt.tolist()
    assert_same_weights(pop.numpy(), ref, "run_net")
    gt = out["traj"].cpu().numpy()
    for i in range(0, n, 53):
        assert_same_weights(gt[:steps[i], i], traj[:steps[i], i])
    c2, n2, flags = pop.classify(1e-4, want_flags=True)
    assert np.array_equal(c2.cpu().numpy(), cls) and np.array_equal(flags.cpu().numpy(), oracle.predicates(do, ref, 1e-4))
    # attack pairs
    pop = eng.Population.from_numpy(dg, w)
    att, vic = np.arange(0, n // 2), np.arange(n // 2, n)
    pop.attack(att, vic)
    ref = w.copy()
    oracle.apply(do, ref, att, ref, vic, ref, vic, n // 2)
    assert_same_weights(pop.numpy(), ref, "attack")
    # forward
    io = 4 if key[0] == 0 else (key[3] if key[0] == 1 else dg.n_weights)
    x = np.random.default_rng(1).uniform(-1, 1, (n, io)).astype(np.float32)
    assert_same_weights(eng.Population.from_numpy(dg, w).forward(x).cpu().numpy(), oracle.forward(do, w, x), "forward")
    # train / learn_from
    for pm in (0, 2):
        pop = eng.Population.from_numpy(dg, w)
        loss = pop.train(4, perm_mode=pm, seed=5, slot0=7, epoch0=1)
        ref = w.copy()
        rl = oracle.train(do, ref, 4, perm_mode=pm, seed=5, slot0=7, epoch0=1)
        assert_same_weights(pop.numpy(), ref, "train perm %d" % pm)
        gl = loss.cpu().numpy()
        assert np.array_equal(np.isnan(gl), np.isnan(rl)) and np.array_equal(gl[~np.isnan(gl)], rl[~np.isnan(rl)])
    t = glorot_population(key, n, seed=15)
    pop, tea = eng.Population.from_numpy(dg, w), eng.Population.from_numpy(dg, t)
    pop.train(3, perm_mode=2, seed=9, teacher=tea, teacher_idx=np.arange(n)[::-1].copy())
    ref = w.copy()
    oracle.train(do, ref, 3, perm_mode=2, seed=9, teacher_w=t, teacher_idx=np.arange(n)[::-1].copy())
    assert_same_weights(pop.numpy(), ref, "learn_from")


@pytest.mark.parametrize("key", [(0, 3, 3, 0), (1, 3, 2, 6), (2, 3, 2, 0)])
def test_generic_shape_population_soup_and_kfv(eng, key):
    """Population-scale soups and the known-fixpoint sweep for shapes outside SRNN_SHAPES: the generic
    generation kernel equals the oracle bit for bit -- one owner, two uneven owners, the peer-pointer entry, the
    multi-generation run behind PopulationSoup -- and so does the generic KFV kernel."""
    import torch
    import soup as soup_mod
    dg, do = both(eng, key)
    assert eng.shape_supported(dg) == 2
    n = 1500
    kw = dict(attacking_rate=0.3, learn_from_rate=0.2, train=1, learn_from_severity=2, remove_divergent=True,
              remove_zero=True, epsilon=1e-4, perm_mode=2)
    pg, po = eng.make_soup_params(**kw), oracle.soup_params(**kw)
    w = glorot_population(key, n, seed=52)
    snap, ref = torch.from_numpy(w).cuda(), w.copy()
    scratch = torch.zeros(eng.soup_scratch_len(n, n), dtype=torch.int32, device="cuda")
    for gen in range(1, 4):
        out = torch.zeros_like(snap)
        action = torch.zeros(n, dtype=torch.uint8, device="cuda")
        loss = torch.zeros(n, dtype=torch.float64, device="cuda")
        eng.soup_sync(dg, pg, snap, n, out, 0, n, 23, gen, scratch, action=action, loss=loss)
        rout = np.zeros_like(ref)
        ra, rl, _ = oracle.soup_sync(do, po, ref, n, rout, 0, n, 23, gen)
        assert_same_weights(out.cpu().numpy(), rout, "gen %d" % gen)
        assert np.array_equal(action.cpu().numpy(), ra)
        gl = loss.cpu().numpy()
        assert np.array_equal(np.isnan(gl), np.isnan(rl)) and np.array_equal(gl[~np.isnan(gl)], rl[~np.isnan(rl)])
        half = n // 2 + 11
        o2 = torch.zeros_like(snap)
        eng.soup_sync(dg, pg, snap, n, o2[:half], 0, half, 23, gen, scratch)
        eng.soup_sync(dg, pg, snap, n, o2[half:], half, n - half, 23, gen, scratch)
        assert torch.equal(o2.view(torch.int32), out.view(torch.int32))
        sh = [snap[:n // 2].clone(), snap[n // 2:].clone()]
        o3 = torch.zeros_like(snap)
        for r in range(2):
            eng.soup_sync_peers(dg, pg, [t.data_ptr() for t in sh], n // 2, r, sh[0].shape[1],
                                o3[r * (n // 2):(r + 1) * (n // 2)], 23, gen, scratch)
        assert torch.equal(o3.view(torch.int32), out.view(torch.int32))
        snap, ref = out, rout
    # PopulationSoup: the multi-generation entry falls back to one generic generation per call
    skw = {k: v for k, v in kw.items() if k != "perm_mode"}
    s = soup_mod.PopulationSoup(key[0], 900, width=key[1], depth=key[2], aggregates=key[3], seed=3, **skw).seed_particles()
    ref = s.local.numpy().copy()
    s.evolve(3)
    for gen in range(1, 4):
        rout = np.zeros_like(ref)
        oracle.soup_sync(do, po, ref, 900, rout, 0, 900, 3, gen)
        ref = rout
    assert_same_weights(s.local.numpy(), ref, "PopulationSoup.evolve(3)")
    assert sum(s.count().values()) == 900
    # known-fixpoint variation around the identity-on-the-first-input fixpoint of a weightwise net of this shape
    if key[0] == 0:
        base = np.zeros(do.n_weights, np.float32)
        off = 0
        for (r, c) in eng.layer_shapes(dg):
            base[off] = 1.0
            off += r * c
        for level, scale in ((3, 1e-3), (6, 1e-6)):
            got = eng.kfv(dg, base, 4000, scale, level, 100, 1e-4, seed=12, want_weights=True)
            ys, zs, wout = oracle.kfv(do, base, 4000, scale, level, 100, 1e-4, seed=12, want_weights=True)
            assert np.array_equal(got["ys"].cpu().numpy(), ys) and np.array_equal(got["zs"].cpu().numpy(), zs)
            assert_same_weights(got["w"].cpu().numpy(), wout, "kfv level %d" % level)
            h = got["hist"].cpu().numpy()
            assert h[0].sum() == 4000 and (h[0] * np.arange(101)).sum() == ys.sum()


def test_generic_shape_facade(eng):
    """The class API accepts any shape: a WeightwiseNeuralNetwork(width=3, depth=3) self-attacks, trains and sits in a
    Soup (object-wise evolve, since soups' fused kernels exist for the instantiated shapes only)."""
    import network
    import soup as soup_mod
    net = network.TrainingNeuralNetworkDecorator(network.ParticleDecorator(network.WeightwiseNeuralNetwork(3, 3)))
    w0 = net.get_weights_flat()
    do = oracle.desc(0, 3, 3, 0)
    net.self_attack(2)
    ref = w0[None].copy()
    oracle.self_apply(do, ref, 2)
    assert_same_weights(net.get_weights_flat(), ref[0])
    assert isinstance(net.train(store_states=False), float)
    s = soup_mod.Soup(6, lambda: network.TrainingNeuralNetworkDecorator(network.WeightwiseNeuralNetwork(3, 3))
                      .with_params(epsilon=1e-4)).with_params(train=1, remove_divergent=True, remove_zero=True)
    s.seed()
    s.evolve(3)
    assert sum(s.count().values()) == 6 and s.time == 3


# ---------------------------------------------------------------- exact-order generations at scale
@pytest.mark.parametrize("key", BENCH_SHAPES)
def test_soup_exact_order_wavefront(eng, key):
    """srnn_soup_exact == the sequential sweep of Soup.evolve (oracle soup_ordered), bit for bit, on 20k particles:
    Philox-drawn pairings (the function the oracle's soup_draw evaluates) and host-supplied pairings."""
    import torch
    dg, do = both(eng, key)
    n = 20000
    kw = dict(attacking_rate=0.3, learn_from_rate=0.2, train=1, learn_from_severity=2, remove_divergent=True,
              remove_zero=True, epsilon=1e-4, perm_mode=2)
    pg, po = eng.make_soup_params(**kw), oracle.soup_params(**kw)
    w = glorot_population(key, n, seed=61)
    arena = torch.from_numpy(w).cuda()
    ref = w.copy()
    scratch = torch.zeros(eng.soup_exact_scratch_len(n), dtype=torch.int32, device="cuda")
    action = torch.zeros(n, dtype=torch.uint8, device="cuda")
    loss = torch.zeros(n, dtype=torch.float64, device="cuda")
    rng = np.random.default_rng(3)
    for gen in range(1, 4):
        if gen < 3:      # device draws
            att, lrn = oracle.soup_draw(po, 29, gen, n)
            rounds = eng.soup_exact(dg, pg, arena, 29, gen, scratch, action=action, loss=loss)
        else:            # host-supplied pairings, including self-attacks and self-teaching
            att = np.where(rng.random(n) < 0.3, rng.integers(0, n, n), -1)
            lrn = np.where(rng.random(n) < 0.2, rng.integers(0, n, n), -1)
            att[:50] = np.arange(50)
            lrn[50:100] = np.arange(50, 100)
            rounds = eng.soup_exact(dg, pg, arena, 29, gen, scratch, att=torch.from_numpy(att).cuda(),
                                    lrn=torch.from_numpy(lrn).cuda(), action=action, loss=loss)
        ra, rl = oracle.soup_ordered(do, po, ref, att, lrn, seed=29, gen=gen)
        assert 3 <= rounds < 200
        assert_same_weights(arena.cpu().numpy(), ref, "generation %d (%d rounds)" % (gen, rounds))
        assert np.array_equal(action.cpu().numpy(), ra)
        gl = loss.cpu().numpy()
        assert np.array_equal(np.isnan(gl), np.isnan(rl)) and np.array_equal(gl[~np.isnan(gl)], rl[~np.isnan(rl)])


def test_soup_exact_order_large_split_launches(eng):
    """From 2^22 particles on an exact-order generation is memsets + an ordinary list-building launch + the cooperative
    kernel (per generation) instead of one cooperative launch for everything: two generations in one srnn_soup_exact_run
    call and one with host-supplied pairings through srnn_soup_exact, against the oracle's sequential sweep."""
    import torch
    dg, do = both(eng, (0, 2, 2, 0))
    n = (1 << 22) + 333
    kw = dict(attacking_rate=0.15, learn_from_rate=0.1, train=0, learn_from_severity=1, remove_divergent=True,
              remove_zero=True, epsilon=1e-4, perm_mode=2)
    pg, po = eng.make_soup_params(**kw), oracle.soup_params(**kw)
    w = glorot_population((0, 2, 2, 0), n, seed=62)
    arena = torch.from_numpy(w).cuda()
    ref = w.copy()
    scratch = torch.full((eng.soup_exact_scratch_len(n),), 0x3C3C3C3C, dtype=torch.int32, device="cuda")
    action = torch.zeros(n, dtype=torch.uint8, device="cuda")
    rounds = torch.zeros(2, dtype=torch.int32, device="cuda")
    eng.soup_exact_run(dg, pg, arena, 41, 1, 2, scratch, action=action, rounds=rounds)
    for gen in (1, 2):
        att, lrn = oracle.soup_draw(po, 41, gen, n)
        ra, _ = oracle.soup_ordered(do, po, ref, att, lrn, seed=41, gen=gen)
    assert_same_weights(arena.cpu().numpy(), ref, "two generations in one call")
    assert np.array_equal(action.cpu().numpy(), ra) and (rounds.cpu().numpy() >= 3).all()
    rng = np.random.default_rng(5)
    att = np.where(rng.random(n) < 0.15, rng.integers(0, n, n), -1)
    lrn = np.where(rng.random(n) < 0.1, rng.integers(0, n, n), -1)
    r3 = eng.soup_exact(dg, pg, arena, 41, 3, scratch, att=torch.from_numpy(att).cuda(), lrn=torch.from_numpy(lrn).cuda(),
                        action=action)
    ra, _ = oracle.soup_ordered(do, po, ref, att, lrn, seed=41, gen=3)
    assert r3 >= 3
    assert_same_weights(arena.cpu().numpy(), ref, "host-supplied pairings")
    assert np.array_equal(action.cpu().numpy(), ra)


def test_population_soup_exact_mode(eng):
    import soup as soup_mod
    n = 6000
    kw = dict(seed=13, attacking_rate=0.2, learn_from_rate=0.2, train=1, remove_divergent=True, remove_zero=True)
    s = soup_mod.PopulationSoup(eng.KIND_WW, n, order="exact", **kw).seed_particles()
    ref = s.local.numpy().copy()
    po = oracle.soup_params(attacking_rate=0.2, learn_from_rate=0.2, train=1, learn_from_severity=1, remove_divergent=True,
                            remove_zero=True, epsilon=1e-4, perm_mode=2)
    for gen in range(1, 4):
        s.evolve()
        att, lrn = oracle.soup_draw(po, 13, gen, n)
        oracle.soup_ordered(oracle.ww(), po, ref, att, lrn, seed=13, gen=gen)
        assert_same_weights(s.local.numpy(), ref, "generation %d" % gen)
    assert s.rounds >= 2 and sum(s.count().values()) == n


def test_soup_exact_many_generations_one_launch(eng):
    """srnn_soup_exact_run: 6 generations in ONE cooperative launch == 6 sequential sweeps of the oracle, bit for bit;
    the per-generation dependency depths land in device memory, nobody synchronises in between."""
    import torch
    dg, do = both(eng, (0, 2, 2, 0))
    n = 30000
    kw = dict(attacking_rate=0.2, learn_from_rate=0.1, train=0, learn_from_severity=1, remove_divergent=True,
              remove_zero=True, epsilon=1e-4, perm_mode=2)
    pg, po = eng.make_soup_params(**kw), oracle.soup_params(**kw)
    w = glorot_population((0, 2, 2, 0), n, seed=7)
    arena = torch.from_numpy(w).cuda()
    ref = w.copy()
    scratch = torch.zeros(eng.soup_exact_scratch_len(n), dtype=torch.int32, device="cuda")
    action = torch.zeros(n, dtype=torch.uint8, device="cuda")
    dead = torch.zeros(n, dtype=torch.uint8, device="cuda")
    rounds = torch.zeros(6, dtype=torch.int32, device="cuda")
    eng.soup_exact_run(dg, pg, arena, 41, 1, 6, scratch, action=action, dead=dead, rounds=rounds)
    for gen in range(1, 7):
        att, lrn = oracle.soup_draw(po, 41, gen, n)
        ra, _ = oracle.soup_ordered(do, po, ref, att, lrn, seed=41, gen=gen)
    assert_same_weights(arena.cpu().numpy(), ref, "after 6 generations")
    assert np.array_equal(action.cpu().numpy(), ra)
    r = rounds.cpu().numpy()
    assert (r >= 3).all() and (r < 100).all(), r


def test_full_size_soup_properties(eng):
    """A 2^22-particle soup generation (the 10 M-particle regime): (i) one owner == three uneven owners == the multi-pass
    path, bit for bit; (ii) two slices of the result equal the oracle's computation of those slices from the full
    snapshot; (iii) every particle is accounted for in the action histogram."""
    import torch
    dg, do = both(eng, (0, 2, 2, 0))
    n = 1 << 22
    kw = dict(attacking_rate=0.1, learn_from_rate=0.1, train=0, learn_from_severity=1, remove_divergent=True,
              remove_zero=True, epsilon=1e-4, perm_mode=2)
    pop = eng.Population(dg, n).init(seed=77)
    pop.self_apply(2)                                    # some nets already tiny / huge, so culling happens
    snap = pop.w
    scratch = torch.zeros(eng.soup_scratch_len(n, n), dtype=torch.int32, device="cuda")
    out = torch.zeros_like(snap)
    action = torch.zeros(n, dtype=torch.uint8, device="cuda")
    eng.soup_sync(dg, eng.make_soup_params(**kw), snap, n, out, 0, n, 5, 1, scratch, action=action)
    # (i) sharding and path invariance
    cuts = [0, n // 3 + 11, 2 * n // 3 - 5, n]
    o2 = torch.zeros_like(snap)
    for a, b in zip(cuts[:-1], cuts[1:]):
        eng.soup_sync(dg, eng.make_soup_params(**kw), snap, n, o2[a:b], a, b - a, 5, 1, scratch)
    assert torch.equal(o2.view(torch.int32), out.view(torch.int32))
    o3 = torch.zeros_like(snap)
    eng.soup_sync(dg, eng.make_soup_params(engine_path=2, **kw), snap, n, o3, 0, n, 5, 1, scratch)
    assert torch.equal(o3.view(torch.int32), out.view(torch.int32))
    # (ii) oracle on two slices of the full snapshot
    snap_h = snap.cpu().numpy()
    po = oracle.soup_params(**kw)
    for slot0 in (0, n - 5000):
        ref = np.zeros((5000, 14), np.float32)
        ra, _, _ = oracle.soup_sync(do, po, snap_h, n, ref, slot0, 5000, 5, 1)
        assert_same_weights(out[slot0:slot0 + 5000].cpu().numpy(), ref, "slice at %d" % slot0)
        assert np.array_equal(action[slot0:slot0 + 5000].cpu().numpy(), ra)
    # (iii) action histogram
    hist = torch.bincount(action.to(torch.int64), minlength=6).cpu().numpy()
    assert hist.sum() == n and hist[3] == 0 and 0.05 * n < hist[1] + hist[2] < 0.3 * n

