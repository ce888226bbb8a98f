"""soup -- drop-in for the reference's code/soup.py on top of the B200 engine.

`Soup` keeps the reference's surface (size, generator, params, particles, historical_particles, time; seed(),
evolve(), count(), with_params(), generate_particle(), get_particle(), print_all(), without_particles()) and
the reference's semantics -- particles visited in index order, every write visible to later particles, RNG
draws taken from Python's `random` in the reference's order (attack coin, [target], learn coin, [target]),
the `action` precedence and the "zero test looks at the old particle" quirk (code/soup.py:51-87).  A whole
generation is ONE launch of the exact-order kernel (srnn_soup_ordered); the particle objects are rows of one
contiguous device arena owned by the soup.

`PopulationSoup` is the population-scale mode (100k - 10M particles, optionally sharded over the GPUs of a box
with one all-gather per generation): no Python object per particle, snapshot-synchronous generations
(srnn_soup_sync) with pairings drawn on the device.  It is a documented departure from the reference's
sequential sweep.
"""
import random

import numpy as np

from network import *  # noqa: F401,F403
from network import _base, _EngineState
import srnn_engine as _eng


def prng():
    return random.random()


class Soup(object):

    def __init__(self, size, generator, **kwargs):
        self.size = size
        self.generator = generator
        self.particles = []
        self.historical_particles = {}
        self.params = dict(attacking_rate=0.1, learn_from_rate=0.1, train=0, learn_from_severity=1)
        self.params.update(kwargs)
        self.time = 0
        self._arena = None
        self._spares = []

    def __copy__(self):
        twin = Soup(self.size, self.generator, **self.params)
        twin.__dict__ = {k: v for k, v in self.__dict__.items()
                         if k not in ('particles', 'historical_particles', '_arena', '_spares')}
        return twin

    def without_particles(self):
        twin = copy.copy(self)  # noqa: F405
        twin.historical_particles = {uid: p.states for uid, p in self.historical_particles.items()}
        return twin

    def with_params(self, **kwargs):
        self.params.update(kwargs)
        return self

    def generate_particle(self):
        particle = ParticleDecorator(self.generator())  # noqa: F405
        self.historical_particles[particle.get_uid()] = particle
        return particle

    def get_particle(self, uid, otherwise=None):
        return self.historical_particles.get(uid, otherwise)

    def seed(self):
        self.particles = [self.generate_particle() for _ in range(self.size)]
        self._arena = None
        return self

    # ------------------------------------------------------------------------------------------- evolve
    def _adopt(self):
        """Move the particles' records into one contiguous arena (row i = particles[i]) if they are engine nets of
        one shape; returns False when the soup is heterogeneous or uses user-overridden operators."""
        bases = [_base(p) for p in self.particles]
        if not bases or any(b is None or not b._fused() or not b._fast_shape for b in bases):
            return False
        if len({b._desc.key() for b in bases}) != 1:
            return False
        if len({float(b.get_params().get('epsilon')) for b in bases}) != 1:
            return False      # the fused sweep tests every particle against one epsilon; the reference uses each particle's own
        if self._arena is None or self._arena.n != len(bases):
            self._arena = _eng.Population(bases[0]._desc, len(bases))
        for i, b in enumerate(bases):
            row = self._arena.w[i:i + 1]
            if b._pop.w.data_ptr() != row.data_ptr():
                b._rebind(_eng.Population(b._desc, 1, device=self._arena.device, ld=self._arena.ld, w=row))
        return True

    def _detach(self, particle):
        """A replaced particle keeps living in historical_particles: give it a private copy of its record."""
        b = _base(particle)
        keep = _eng.Population(b._desc, 1)
        keep.w.copy_(b._pop.w)
        b._pop = keep
        b._touch()

    def _top_up_spares(self, want):
        """Fresh nets for respawns, produced by the user's generator() ahead of time (the kernel consumes them
        in particle order).  They become particles -- and get their uid -- only when consumed."""
        while len(self._spares) < want:
            self._spares.append(self.generator())

    def evolve(self, iterations=1):
        for _ in range(iterations):
            self.time += 1
            if self._adopt():
                self._evolve_generation_fused()
            else:
                self._evolve_generation_objectwise()

    def _evolve_generation_fused(self):
        import torch
        n = len(self.particles)
        p = self.params
        att, lrn = np.full(n, -1, np.int64), np.full(n, -1, np.int64)
        for i in range(n):                                   # the reference's draw order, code/soup.py:56-63
            if prng() < p.get('attacking_rate'):
                att[i] = int(prng() * n)
            if prng() < p.get('learn_from_rate'):
                lrn[i] = int(prng() * n)
        culls = bool(p.get('remove_divergent')) or bool(p.get('remove_zero'))
        spares = None
        if culls:
            self._top_up_spares(n)
            spares = torch.cat([_base(s)._pop.w for s in self._spares[:n]], dim=0).contiguous()
        eps = _base(self.particles[0]).get_params().get('epsilon')
        sp = _eng.make_soup_params(attacking_rate=p.get('attacking_rate'), learn_from_rate=p.get('learn_from_rate'),
                                   train=p.get('train', 0), learn_from_severity=p.get('learn_from_severity', 1),
                                   remove_divergent=p.get('remove_divergent'), remove_zero=p.get('remove_zero'),
                                   epsilon=eps, lr=0.01, perm_mode=_eng.PERM_PHILOX)
        out = self._arena.soup_ordered(sp, att, lrn, seed=_EngineState.seed ^ 0x5eed50, gen=self.time, want_logged=True,
                                       spares=spares)
        action, loss, logged = out[0].cpu().numpy(), out[1].cpu().numpy(), out[2].cpu().numpy()
        for b in (_base(q) for q in self.particles):
            b._touch()
        # replay the bookkeeping of code/soup.py:55-87 in particle order (uids of respawned slots change mid-sweep)
        train = p.get('train', 0)
        for i in range(n):
            particle = self.particles[i]
            desc = {'time': self.time}
            if att[i] >= 0:
                desc['action'] = 'attacking'
                desc['counterpart'] = self.particles[att[i]].get_uid()
            if lrn[i] >= 0:
                desc['action'] = 'learn_from'
                desc['counterpart'] = self.particles[lrn[i]].get_uid()
            if train > 0:
                desc.update(fitted=train, loss=float(loss[i]), action='train_self', counterpart=None)
            if action[i] in (4, 5):                          # divergent_dead / zweo_dead: slot i was respawned
                fresh_net = self._spares.pop(0)
                self._detach(particle)
                _base(particle)._pop.w[0, :logged.shape[1]].copy_(torch.from_numpy(logged[i]))
                fresh = ParticleDecorator(fresh_net)  # noqa: F405
                self.historical_particles[fresh.get_uid()] = fresh
                row = self._arena.w[i:i + 1]
                fb = _base(fresh)
                fb._release()
                fb._pop = _eng.Population(fb._desc, 1, device=self._arena.device, ld=self._arena.ld, w=row)
                fb._touch()      # later attackers of this sweep may already have rewritten the row: drop cached predicates
                self.particles[i] = fresh
                desc['action'] = 'divergent_dead' if action[i] == 4 else 'zweo_dead'
                desc['counterpart'] = fresh.get_uid()
            particle._log(particle._state_of(logged[i, :_base(particle)._desc.n_weights].copy(), **desc))

    def _evolve_generation_objectwise(self):
        """Heterogeneous soups / user-overridden operators: the reference's loop on per-object engine calls."""
        p = self.params
        for particle_id, particle in enumerate(self.particles):
            desc = {'time': self.time}
            if prng() < p.get('attacking_rate'):
                other = self.particles[int(prng() * len(self.particles))]
                particle.attack(other)
                desc['action'], desc['counterpart'] = 'attacking', other.get_uid()
            if prng() < p.get('learn_from_rate'):
                other = self.particles[int(prng() * len(self.particles))]
                for _ in range(p.get('learn_from_severity', 1)):
                    particle.learn_from(other)
                desc['action'], desc['counterpart'] = 'learn_from', other.get_uid()
            for _ in range(p.get('train', 0)):
                particle.compiled()
                loss = particle.train(store_states=False)
                desc.update(fitted=p.get('train', 0), loss=loss, action='train_self', counterpart=None)
            if p.get('remove_divergent') and particle.is_diverged():
                fresh = self.generate_particle()
                self.particles[particle_id] = fresh
                desc['action'], desc['counterpart'] = 'divergent_dead', fresh.get_uid()
            if p.get('remove_zero') and particle.is_zero():
                fresh = self.generate_particle()
                self.particles[particle_id] = fresh
                desc['action'], desc['counterpart'] = 'zweo_dead', fresh.get_uid()
            particle.save_state(**desc)

    # ------------------------------------------------------------------------------------------- count
    def count(self):
        counters = dict(divergent=0, fix_zero=0, fix_other=0, fix_sec=0, other=0)
        if self.particles and self._adopt():
            eps = _base(self.particles[0]).get_params().get('epsilon')
            _, cnt = self._arena.classify(eps, with_fix_sec=True)
            for name, c in zip(_eng.CLS_NAMES, cnt.cpu().numpy().tolist()):
                counters[name] = int(c)
            return counters
        for particle in self.particles:
            if particle.is_diverged():
                counters['divergent'] += 1
            elif particle.is_fixpoint():
                counters['fix_zero' if particle.is_zero() else 'fix_other'] += 1
            elif particle.is_fixpoint(2):
                counters['fix_sec'] += 1
            else:
                counters['other'] += 1
        return counters

    def print_all(self):
        for particle in self.particles:
            particle.print_weights()
            print(particle.is_fixpoint())


class _CudaBackend:
    """PopulationSoup's compute backend: the B200 engine.  (Tests inject a CPU stand-in to exercise the sharding
    logic under gloo without a GPU; the product never uses anything else.)"""

    def __init__(self, device=None):
        import torch
        _eng._torch()
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)

    def alloc(self, shape, dtype):
        import torch
        return torch.zeros(shape, dtype=dtype, device=self.device)

    def init(self, desc, w, seed, slot0, generation):
        _eng.Population(desc, w.shape[0], device=self.device, ld=w.shape[1], w=w).init(seed, slot0=slot0, generation=generation)

    def soup_sync(self, desc, sp_kwargs, snap, n_total, w_out, slot0, n, seed, gen, scratch, action, dead, loss=None):
        _eng.soup_sync(desc, _eng.make_soup_params(**sp_kwargs), snap, n_total, w_out, slot0, n, seed, gen, scratch,
                       action=action, dead=dead, loss=loss)

    def soup_sync_peers(self, desc, sp_kwargs, shard_ptrs, shard_len, my_shard, ld, w_out, seed, gen, scratch, action,
                        dead, stream=None, loss=None):
        sp = sp_kwargs if isinstance(sp_kwargs, _eng.SoupParams) else _eng.make_soup_params(**sp_kwargs)
        _eng.soup_sync_peers(desc, sp, shard_ptrs, shard_len, my_shard, ld, w_out, seed, gen, scratch, action=action,
                             dead=dead, stream=stream, loss=loss)

    def peer_arena(self, n_local, nw, group):
        """Two ping-pong arenas in NVLink-mapped symmetric memory: returns (tensor (2, n_local, nw), handle) or None
        when peer memory cannot be set up (then the soup falls back to the all-gather exchange)."""
        try:
            import torch
            import torch.distributed as dist
            import torch.distributed._symmetric_memory as symm
            buf = symm.empty((2, n_local, nw), dtype=torch.float32, device=self.device)
            hdl = symm.rendezvous(buf, group if group is not None else dist.group.WORLD)
            if len(hdl.buffer_ptrs) != hdl.world_size or any(int(q) == 0 for q in hdl.buffer_ptrs):
                return None
            buf.zero_()
            return buf, hdl
        except Exception as exc:   # no P2P / fabric support on this box
            import warnings
            warnings.warn("PopulationSoup: peer memory unavailable (%s); using the all-gather exchange" % (exc,))
            return None

    def peer_lists(self, n_local, n_total, group):
        """Attack-list arrays in peer-mapped memory: (2 parities) x (head[n_local] | next[n_total]) int32 per rank,
        followed by 64 words for the generation flags of the fused cross-GPU signal; returns (tensor, handle) or None."""
        try:
            import torch
            import torch.distributed as dist
            import torch.distributed._symmetric_memory as symm
            buf = symm.empty((2 * (n_local + n_total) + 64,), dtype=torch.int32, device=self.device)
            hdl = symm.rendezvous(buf, group if group is not None else dist.group.WORLD)
            if len(hdl.buffer_ptrs) != hdl.world_size or any(int(q) == 0 for q in hdl.buffer_ptrs):
                return None
            buf.fill_(-1)
            return buf, hdl
        except Exception:
            return None

    def soup_push_lists(self, sp, n_shards, shard_len, my_shard, seed, gen, heads, nexts, stream=None):
        _eng.soup_push_lists(sp, n_shards, shard_len, my_shard, seed, gen, heads, nexts, stream=stream)

    def soup_chain_peers(self, desc, sp, shard_ptrs, shard_len, my_shard, ld, w_out, seed, gen, heads_cur, nexts_cur,
                         heads_nxt, nexts_nxt, action, dead, stream=None, flags=None, wait_gen=0, loss=None, scratch=None):
        _eng.soup_chain_peers(desc, sp, shard_ptrs, shard_len, my_shard, ld, w_out, seed, gen, heads_cur, nexts_cur,
                              heads_nxt, nexts_nxt, action=action, dead=dead, stream=stream, flags=flags, wait_gen=wait_gen,
                              loss=loss, scratch=scratch)

    def soup_exact(self, desc, sp_kwargs, w, seed, gen0, n_gens, scratch, action, dead, rounds, loss=None):
        sp = _eng.make_soup_params(**sp_kwargs)
        if loss is not None and n_gens == 1:       # the single-generation entry also returns the losses
            _eng.soup_exact(desc, sp, w, seed, gen0, scratch, action=action, loss=loss, dead=dead, want_rounds=False)
            return
        _eng.soup_exact_run(desc, sp, w, seed, gen0, n_gens, scratch, action=action, dead=dead, rounds=rounds)

    def soup_log_watched(self, desc, sp_kwargs, *args):
        _eng.soup_log_watched(desc, _eng.make_soup_params(**sp_kwargs), *args)

    def soup_sync_run(self, desc, sp_kwargs, w_a, w_b, seed, gen0, n_gens, scratch, action, dead):
        _eng.soup_sync_run(desc, _eng.make_soup_params(**sp_kwargs), w_a, w_b, seed, gen0, n_gens, scratch,
                           action=action, dead=dead)

    def classify(self, desc, w, eps, with_fix_sec):
        return _eng.Population(desc, w.shape[0], device=self.device, ld=w.shape[1], w=w).classify(eps, with_fix_sec)[1]

    def assign_uids(self, dead, uid, base, tmp, n_dead):
        _eng.assign_uids(dead, uid, base, tmp, n_dead)


def shard_of(size, world, rank):
    """Contiguous block of particle indices owned by `rank`: (slot0, n_local).  Global index = slot0 + local."""
    if size % world:
        raise ValueError("population size must be a multiple of the number of ranks")
    n_local = size // world
    return rank * n_local, n_local


def uid_base_of(rank, dead_counts, next_uid):
    """First uid handed out on `rank` this generation and the new global counter: respawned particles get uids in
    global particle order (= the order generate_particle would be called, code/soup.py:78,83)."""
    return next_uid + int(sum(dead_counts[:rank])), next_uid + int(sum(dead_counts))


_SIDE_STREAMS = {}


def evolve_together(soups, iterations=1):
    """Advance several PopulationSoups in lock step -- the kinds of a mixed soup (separate arenas, partners within
    kind).  Results equal `for s in soups: s.evolve(iterations)` bit for bit: the soups do not
    interact.  For sharded soups with the peer-memory exchange every kind runs its own chain of generations and
    cross-GPU barriers on its own stream (they are small and latency-bound at 8 GPUs: three of them fill the SMs
    together, and nobody waits for another kind's barrier); anything else falls back to the plain loop."""
    soups = list(soups)
    fast = len(soups) > 1 and all(isinstance(s, PopulationSoup) and s._peer is not None and s.order == "snapshot"
                                  for s in soups) and len({s._lists is None for s in soups}) == 1
    if not fast:
        for s in soups:
            s.evolve(iterations)
        return soups
    import torch
    dev = soups[0].w.device
    side = _SIDE_STREAMS.setdefault((dev.index, len(soups)), [torch.cuda.Stream(device=dev) for _ in soups])
    main = torch.cuda.current_stream(dev)
    sps = [_eng.make_soup_params(**s._sp()) if isinstance(s.backend, _CudaBackend) else s._sp() for s in soups]
    barrier = soups[0]._peer[1].barrier
    if iterations > 0 and any(s._chain_stale() for s in soups):
        barrier()                         # nobody is still pushing into anybody's lists
        for s in soups:
            s._chain_reset()
        barrier()                         # every rank's heads are reset before anybody pushes
        for s, sp in zip(soups, sps):
            s._chain_push_first(sp)
        barrier()                         # the first generation's lists are complete on every rank
    # Each kind runs its own chain  generation -> cross-GPU barrier -> generation ...  on its own stream, with its own
    # symmetric-memory handle: the kinds never wait for each other, so one kind's barrier (a round trip over NVLink)
    # travels under the other kinds' kernels instead of idling the GPU once per generation.
    fork = torch.cuda.Event()
    fork.record(main)
    for st in side:
        st.wait_event(fork)
    for _ in range(iterations):           # enqueued generation by generation so that all chains start at once
        for s, sp, st in zip(soups, sps, side):
            if s._lists is not None and not s.track_uids and s._watch is None:
                s.time += 1               # chained exchange: one C call on an explicit stream, nothing else to order
                s._peer_generation(sp, stream=st)
                s._peer_flip()
                continue
            with torch.cuda.stream(st):
                s.time += 1
                if not s._peer_generation(sp, stream=st):
                    s._peer[1].barrier()  # all ranks finished this kind's generation g before anyone overwrites it
                s._peer_flip()
    for st in side:
        join = torch.cuda.Event()
        join.record(st)
        main.wait_event(join)
    return soups


class PopulationSoup:
    """Population-scale soup of ONE net shape, device-resident, optionally sharded by particle index over the ranks
    of a torch.distributed process group (one rank per GPU).  One generation =
        [all-gather of the record arena over NVLink when sharded] -> srnn_soup_sync on the local slice
    Pairings, SGD sample orders and respawns are Philox functions of (seed, global particle index, generation),
    so the result does not depend on how the population is sharded."""

    def __init__(self, kind, size, width=2, depth=2, aggregates=0, seed=0, group=None, device=None, backend=None,
                 exchange="auto", order="snapshot", sharded=True, **params):
        import torch
        import torch.distributed as dist
        self.desc = _eng.make_desc(kind, width, depth, aggregates)
        self.group = group
        # sharded=False: this process owns the whole population even inside a process group (the single-owner twin the
        # sharded result is checked against)
        self.world = dist.get_world_size(group) if (sharded and dist.is_available() and dist.is_initialized()) else 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        self.size = int(size)
        self.slot0, self.n_local = shard_of(self.size, self.world, self.rank)
        self.seed, self.time = int(seed), 0
        self.params = dict(attacking_rate=0.1, learn_from_rate=0.1, train=0, learn_from_severity=1,
                           remove_divergent=False, remove_zero=False, epsilon=1e-4, perm_mode=_eng.PERM_PHILOX,
                           engine_path=0)
        self.params.update(params)
        # order="snapshot": partners read from the start-of-generation arena (default, shards over GPUs);
        # order="exact": the reference's sequential sweep, bit for bit, as a parallel wavefront (unsharded soups)
        if order not in ("snapshot", "exact"):
            raise ValueError("order must be 'snapshot' or 'exact'")
        if order == "exact" and self.world > 1:
            raise ValueError("exact-order generations are implemented for unsharded soups")
        self.order, self._rounds = order, None
        self.backend = backend or _CudaBackend(device)
        nw, B = self.desc.n_weights, self.backend
        # exchange of a sharded soup: "peer" = partner records are read straight from the owning GPU's arena over
        # NVLink inside the generation kernel (srnn_soup_sync_peers); "allgather" = NCCL all-gather of the arenas,
        # then srnn_soup_sync; "auto" = peer when symmetric memory can be set up
        self._peer = None
        if self.world > 1 and exchange in ("auto", "peer", "peer_unchained") and hasattr(B, "peer_arena"):
            self._peer = B.peer_arena(self.n_local, nw, group)
            if self._peer is None and exchange != "auto":
                raise RuntimeError("peer-memory exchange requested but symmetric memory is unavailable")
        # "peer": attack lists built by the attackers' ranks, one launch per generation (srnn_soup_chain_peers);
        # "peer_unchained": every rank re-draws all pairings and builds its lists locally (srnn_soup_sync_peers)
        # "auto" picks by measurement (mixed 10 M soup, generations/s, unchained vs chained): 2 GPUs 2020 vs 1792,
        # 4 GPUs 3263 vs 3147, 8 GPUs 5534 vs 5716 -- the chained form wins once the all-particle list pass dominates
        unchained = exchange == "peer_unchained" or (exchange == "auto" and self.world <= 4)
        self.exchange = ("peer_unchained" if unchained else "peer") if self._peer is not None else (
            "allgather" if self.world > 1 else "none")
        self._shard_arrays = None
        self._lists = None          # chained peer exchange: attack lists in peer memory, built by the attackers' ranks
        self._lists_valid = None    # (generation, parameters) the lists of parity generation & 1 were built for
        if self._peer is not None and hasattr(B, "peer_lists") and self.params.get("engine_path", 0) == 0 \
                and self.exchange == "peer":
            self._lists = B.peer_lists(self.n_local, self.size, group)
        if self._peer is not None:
            buf, _ = self._peer
            self._cur = 0
            self.w, self._next, self._snap = buf[0], buf[1], None
        else:
            self.w = B.alloc((self.n_local, nw), torch.float32)
            self._next = B.alloc((self.n_local, nw), torch.float32)
            self._snap = B.alloc((self.size, nw), torch.float32) if self.world > 1 else None
        words = _eng.soup_exact_scratch_len(self.size) if order == "exact" else _eng.soup_scratch_len(self.size, self.n_local)
        self._scratch = B.alloc((words,), torch.int32)
        self.dead = B.alloc((self.n_local,), torch.uint8)
        self.action = B.alloc((self.n_local,), torch.uint8)
        self.uid = B.alloc((self.n_local,), torch.int64)
        self.uid.copy_(torch.arange(self.slot0, self.slot0 + self.n_local, dtype=torch.int64))
        self._uid_tmp = B.alloc(((self.n_local + 1023) // 1024 + 1,), torch.int64)
        self._n_dead = B.alloc((1,), torch.int64)
        self.next_uid = self.size
        self.track_uids = False
        self._watch = None          # (local indices, every, {local index -> list of state dicts})

    @property
    def rounds(self):
        """Depth of the dependency DAG of the last exact-order generation (reads one int back from the device)."""
        return 0 if self._rounds is None else int(self._rounds[-1].item())

    @property
    def local(self):
        """The local slice as an engine Population (CUDA backend only)."""
        return _eng.Population(self.desc, self.n_local, device=self.w.device, ld=self.w.shape[1], w=self.w)

    def with_params(self, **kwargs):
        self.params.update(kwargs)
        return self

    def seed_particles(self):
        """Soup.seed: fresh Keras-style nets for every local particle."""
        self.backend.init(self.desc, self.w, self.seed, self.slot0, 0)
        if self._peer is not None:
            self._peer[1].barrier()          # everybody's arena is initialised before anyone reads a peer
        return self

    def _sp(self):
        sp = {k: self.params[k] for k in ("attacking_rate", "learn_from_rate", "train", "learn_from_severity",
                                          "remove_divergent", "remove_zero", "epsilon", "perm_mode", "engine_path")}
        if self._watch is not None:
            sp["dying"] = self._watch["dying"]      # pre-respawn records of the particles that die (trajectory log)
        return sp

    def evolve(self, iterations=1):
        import torch.distributed as dist
        sp, B = self._sp(), self.backend
        if self.order == "exact":
            # the reference's sequential sweep as a wavefront: whole generations in one cooperative launch, no host sync;
            # uid tracking / watching need the host between generations, so they go one generation per launch
            import torch
            step = 1 if (self.track_uids or self._watch is not None) else iterations
            done = 0
            while done < iterations:
                k = min(step, iterations - done)
                self._rounds = B.alloc((k,), torch.int32) if hasattr(B, "alloc") else None
                self._pre_generation()
                B.soup_exact(self.desc, sp, self.w, self.seed, self.time + 1, k, self._scratch, self.action, self.dead,
                             self._rounds, **self._loss_kw())
                self.time += k
                done += k
                if self.track_uids:
                    self._assign_uids()
                self._log_watched()
            return self
        if self.world == 1 and not self.track_uids and self._watch is None and iterations > 1 and hasattr(B, "soup_sync_run"):
            # unsharded: all generations in one C call (no per-generation host overhead)
            B.soup_sync_run(self.desc, sp, self.w, self._next, self.seed, self.time + 1, iterations, self._scratch,
                            self.action, self.dead)
            self.time += iterations
            if iterations & 1:
                self.w, self._next = self._next, self.w
            return self
        if self._peer is not None and iterations > 0 and self._chain_stale():
            spc = _eng.make_soup_params(**sp) if isinstance(B, _CudaBackend) else sp
            self._peer[1].barrier()
            self._chain_reset()
            self._peer[1].barrier()
            self._chain_push_first(spc)
            self._peer[1].barrier()
        for _ in range(iterations):
            self.time += 1
            if self._peer is not None:
                if not self._peer_generation(sp):
                    self._peer[1].barrier()  # all ranks finished reading generation g before anyone overwrites it
                self._peer_flip()
                continue
            if self.world > 1:
                dist.all_gather_into_tensor(self._snap, self.w, group=self.group)   # NCCL over NVLink on the box
                snap = self._snap
            else:
                snap = self.w
            self._pre_generation()
            B.soup_sync(self.desc, sp, snap, self.size, self._next, self.slot0, self.n_local, self.seed, self.time,
                        self._scratch, self.action, self.dead, **self._loss_kw())
            self.w, self._next = self._next, self.w
            if self.track_uids:
                self._assign_uids()
            self._log_watched()
        return self

    def _list_tables(self):
        """Pointer tables of every rank's list arrays: [parity] -> (heads, nexts); [2] = the flag arrays."""
        if getattr(self, "_list_ptrs", None) is None:
            stride = (self.n_local + self.size) * 4
            self._list_ptrs = []
            for par in (0, 1):
                heads = _eng.shard_ptr_array([int(q) + par * stride for q in self._lists[1].buffer_ptrs])
                nexts = _eng.shard_ptr_array([int(q) + par * stride + self.n_local * 4 for q in self._lists[1].buffer_ptrs])
                self._list_ptrs.append((heads, nexts))
            self._list_ptrs.append(_eng.shard_ptr_array([int(q) + 2 * stride for q in self._lists[1].buffer_ptrs]))
        return self._list_ptrs

    def _params_key(self):
        return tuple(sorted((k, float(v) if isinstance(v, float) else v) for k, v in self._sp().items() if k != "dying"))

    def _chain_stale(self):
        """True when the lists of generation time+1 still have to be built (first run, parameters changed, or the soup
        was advanced some other way).  Every rank takes the same decision.  The caller then runs the bootstrap:
        barrier, _chain_reset, barrier, _chain_push_first, barrier."""
        return self._lists is not None and self._lists_valid != (self.time + 1, self._params_key())

    def _chain_reset(self):
        if self._lists is None or not self._chain_stale():
            return
        per = self.n_local + self.size
        buf = self._lists[0]
        buf[:self.n_local].fill_(-1)
        buf[per:per + self.n_local].fill_(-1)
        flags = buf[2 * per:2 * per + 64]
        flags.fill_(self.time)               # "every rank has completed generation `time`"
        flags[self.world] = 0                # the CTA counter of the fused signal
        self._lists_valid = None

    def _chain_push_first(self, sp):
        if self._lists is None or self._lists_valid is not None:
            return
        heads, nexts = self._list_tables()[(self.time + 1) & 1]
        self.backend.soup_push_lists(sp, self.world, self.n_local, self.rank, self.seed, self.time + 1, heads, nexts)
        self._lists_valid = (self.time + 1, self._params_key())

    def _peer_generation(self, sp, stream=None):
        """Launch generation self.time of a peer-exchange soup on `stream` / the current stream.  Returns True when the
        launch carries its own cross-GPU ordering (chained exchange: generation flags), False when the caller still
        has to barrier before the next generation."""
        if self._shard_arrays is None:      # the peers' two ping-pong arenas, as ready-made pointer tables
            stride = self.n_local * self.desc.n_weights * 4
            self._shard_arrays = [_eng.shard_ptr_array([int(q) + c * stride for q in self._peer[1].buffer_ptrs])
                                  for c in (0, 1)]
        if self._lists is not None and self._lists_valid is not None and self._lists_valid[0] == self.time:
            # chained: one launch; it also pushes generation time+1's attacks onto the owners' lists
            if not isinstance(sp, _eng.SoupParams):
                sp = _eng.make_soup_params(**sp)
            tabs = self._list_tables()
            (hc, nc), (hn, nn) = tabs[self.time & 1], tabs[(self.time + 1) & 1]
            self._pre_generation()
            self.backend.soup_chain_peers(self.desc, sp, self._shard_arrays[self._cur], self.n_local, self.rank,
                                          self.desc.n_weights, self._next, self.seed, self.time, hc, nc, hn, nn,
                                          self.action, self.dead, stream=stream, flags=tabs[2], wait_gen=self.time - 1,
                                          scratch=self._scratch, **self._loss_kw())
            self._lists_valid = (self.time + 1, self._lists_valid[1])
            return True
        self._pre_generation()
        self.backend.soup_sync_peers(self.desc, sp, self._shard_arrays[self._cur], self.n_local, self.rank,
                                     self.desc.n_weights, self._next, self.seed, self.time, self._scratch, self.action,
                                     self.dead, stream=stream, **self._loss_kw())
        return False

    def _peer_flip(self):
        self._cur ^= 1
        self.w, self._next = self._next, self.w
        if self.track_uids:
            self._assign_uids()
        self._log_watched()

    def _assign_uids(self):
        """Deterministic global uids for respawned particles: rank offsets from an all-gather of per-rank counts."""
        import torch
        import torch.distributed as dist
        logging = self._logging_now()
        if logging:
            self._log_before_uids()
        mine = self.dead.sum(dtype=torch.int64).reshape(1)
        if self.world > 1:
            counts = torch.zeros(self.world, dtype=torch.int64, device=mine.device)
            dist.all_gather_into_tensor(counts, mine, group=self.group)
            counts = counts.cpu().tolist()
        else:
            counts = [int(mine.item())]
        base, self.next_uid = uid_base_of(self.rank, counts, self.next_uid)
        self.backend.assign_uids(self.dead, self.uid, base, self._uid_tmp, self._n_dead)
        if logging:
            self._log_after_uids()

    # ---- trajectory logging at scale: the reference logs every particle every generation
    #      (code/soup.py:87); a 10 M-particle soup logs a chosen subset, every k-th generation, in the same schema.
    #      The states are gathered ON THE DEVICE into a ring buffer (srnn_soup_log_watched: weights, action, counterpart,
    #      loss, the record a dying particle had before its respawn) and come to the host once per `capacity` logged
    #      generations -- not with a .cpu() per generation.
    def watch(self, particles, every=1, capacity=64):
        """Record the states of the given GLOBAL particle indices (those owned by this rank) every `every` generations.
        Turns on uid tracking so respawned particles start a new trajectory under a new uid, as in the reference."""
        import torch
        mine = [int(g) - self.slot0 for g in particles if self.slot0 <= int(g) < self.slot0 + self.n_local]
        B, nw, W, K = self.backend, self.desc.n_weights, len(mine), int(capacity)
        idx = B.alloc((max(W, 1),), torch.int32)
        idx[:W] = torch.as_tensor(mine, dtype=torch.int32)
        cap = max(1024, self.n_local // 64)
        self._watch = dict(
            idx=idx[:W], every=int(every), K=K, rows=0, times=[], log={},
            ring_w=B.alloc((K, max(W, 1), nw), torch.float32), ring_dying=B.alloc((K, max(W, 1), nw), torch.float32),
            ring_action=B.alloc((K, max(W, 1)), torch.uint8), ring_uid=B.alloc((K, max(W, 1)), torch.int64),
            ring_cp=B.alloc((K, max(W, 1)), torch.int64), ring_cpuid=B.alloc((K, max(W, 1)), torch.int64),
            ring_newuid=B.alloc((K, max(W, 1)), torch.int64), ring_loss=B.alloc((K, max(W, 1)), torch.float64),
            dying=(B.alloc((cap, nw), torch.float32), B.alloc((cap,), torch.int64), B.alloc((1,), torch.int32)),
            loss=B.alloc((self.n_local,), torch.float64))
        self.track_uids = True
        if W:                                   # ParticleDecorator.__init__ logs the 'init' state at time 0
            w0 = self.w[self._watch["idx"].long(), :nw].cpu().numpy()
            u0 = self.uid[self._watch["idx"].long()].cpu().numpy()
            name = _eng.KIND_NAMES[self.desc.kind]
            for row in range(W):
                if np.isfinite(w0[row]).all():
                    self._watch["log"].setdefault(int(u0[row]), []).append(
                        {'class': name, 'weights': w0[row].copy(), 'time': 0, 'action': 'init', 'counterpart': None})
        return self

    def _logging_now(self):
        return self._watch is not None and self._watch["idx"].numel() > 0 and self.time % self._watch["every"] == 0

    def _pre_generation(self):
        """Called right before generation self.time is launched."""
        if self._watch is not None:
            self._watch["dying"][2].zero_()

    def _loss_kw(self):
        return {} if self._watch is None else {"loss": self._watch["loss"]}

    def _log_before_uids(self):
        """After generation self.time, before the respawned slots get their new uids: one ring row, on the device."""
        import torch
        import torch.distributed as dist
        wt = self._watch
        r = wt["rows"]
        self.backend.soup_log_watched(self.desc, self._sp(), self.w, self.size, self.slot0, wt["idx"], self.seed, self.time,
                                      self.action, wt["loss"], self.uid, wt["ring_w"][r], wt["ring_dying"][r],
                                      wt["ring_action"][r], wt["ring_uid"][r], wt["ring_cp"][r], wt["ring_loss"][r])
        # counterpart: global index -> the uid that particle had when the generation started (other_particle.get_uid())
        uid_all = self.uid
        if self.world > 1:
            uid_all = torch.zeros(self.size, dtype=torch.int64, device=self.uid.device)
            dist.all_gather_into_tensor(uid_all, self.uid, group=self.group)
        W = wt["idx"].numel()
        cp = wt["ring_cp"][r, :W]
        wt["ring_cpuid"][r, :W] = torch.where(cp >= 0, uid_all[cp.clamp(min=0)], cp)

    def _log_after_uids(self):
        wt = self._watch
        r, W = wt["rows"], wt["idx"].numel()
        wt["ring_newuid"][r, :W] = self.uid[wt["idx"].long()]
        wt["times"].append(self.time)
        wt["rows"] = r + 1
        if wt["rows"] == wt["K"]:
            self._drain_watch()

    def _log_watched(self):
        """Kept for the generation loops: the ring row of a generation is written around the uid assignment, see
        _assign_uids; nothing to do here."""
        return

    def _drain_watch(self):
        """Ring -> host: ONE copy per buffer, then the reference's state dicts (code/network.py:185-198, soup.py:55-87)."""
        wt = self._watch
        if wt is None or wt["rows"] == 0:
            return
        r, W = wt["rows"], wt["idx"].numel()
        name = _eng.KIND_NAMES[self.desc.kind]
        w = wt["ring_w"][:r, :W].cpu().numpy()
        dying = wt["ring_dying"][:r, :W].cpu().numpy()
        act = wt["ring_action"][:r, :W].cpu().numpy()
        uid = wt["ring_uid"][:r, :W].cpu().numpy()
        cpuid = wt["ring_cpuid"][:r, :W].cpu().numpy()
        newuid = wt["ring_newuid"][:r, :W].cpu().numpy()
        loss = wt["ring_loss"][:r, :W].cpu().numpy()
        train = int(self.params.get("train", 0) or 0)
        log = wt["log"]
        for row, t in enumerate(wt["times"]):
            for j in range(W):
                a = int(act[row, j])
                state = {'class': name, 'time': t}
                if a in (4, 5):                # the old particle's last state, then a new trajectory under the new uid
                    if train > 0:
                        state.update(fitted=train, loss=float(loss[row, j]))
                    state.update(action=_eng.ACT_NAMES[a], counterpart=int(newuid[row, j]))
                    if np.isfinite(dying[row, j]).all():       # make_state drops non-finite states (network.py:187-188)
                        state['weights'] = dying[row, j].copy()
                        log.setdefault(int(uid[row, j]), []).append(state)
                    if np.isfinite(w[row, j]).all():
                        log.setdefault(int(newuid[row, j]), []).append(
                            {'class': name, 'weights': w[row, j].copy(), 'time': 0, 'action': 'init', 'counterpart': None})
                    continue
                if not np.isfinite(w[row, j]).all():
                    continue
                state['weights'] = w[row, j].copy()
                if a in (1, 2):
                    state.update(action=_eng.ACT_NAMES[a], counterpart=int(cpuid[row, j]))
                elif a == 3:
                    state.update(fitted=train, loss=float(loss[row, j]), action='train_self', counterpart=None)
                log.setdefault(int(uid[row, j]), []).append(state)
        wt["rows"], wt["times"] = 0, []

    def historical_particles(self):
        """uid -> list of state dicts for the watched particles: the structure Soup.without_particles() pickles and the
        reference's visualization.py reads (code/visualization.py:27-40).  Drains the device ring first."""
        if self._watch is None:
            return {}
        self._drain_watch()
        return self._watch["log"]

    # ---- checkpoint / resume (the reference can only dill-dump finished experiments, code/experiment.py:56-59)
    def state_dict(self):
        """Everything needed to continue this rank's shard bit-for-bit: records, uids, clocks, parameters."""
        return dict(kind=self.desc.kind, width=self.desc.width, depth=self.desc.depth, aggregates=self.desc.aggregates,
                    size=self.size, world=self.world, rank=self.rank, seed=self.seed, time=self.time,
                    next_uid=self.next_uid, params=dict(self.params), w=self.w.detach().cpu().clone(),
                    uid=self.uid.detach().cpu().clone())

    def load_state_dict(self, state):
        for key in ("kind", "width", "depth", "size", "world", "rank"):
            have = getattr(self.desc, key) if key in ("kind", "width", "depth") else getattr(self, key)
            if int(state[key]) != int(have):
                raise ValueError("checkpoint %s=%r does not match this soup (%r)" % (key, state[key], have))
        self.seed, self.time, self.next_uid = int(state["seed"]), int(state["time"]), int(state["next_uid"])
        self.params.update(state["params"])
        self.w.copy_(state["w"])
        self.uid.copy_(state["uid"])
        if self._peer is not None:
            self._peer[1].barrier()
        return self

    def save(self, path):
        """One file per rank: `path` gets a .rank{r} suffix when the soup is sharded."""
        import torch
        torch.save(self.state_dict(), path if self.world == 1 else "%s.rank%d" % (path, self.rank))

    def load(self, path):
        import torch
        return self.load_state_dict(torch.load(path if self.world == 1 else "%s.rank%d" % (path, self.rank),
                                               weights_only=False))

    def count(self, with_fix_sec=True):
        """Soup.count over the whole (possibly sharded) population."""
        import torch.distributed as dist
        cnt = self.backend.classify(self.desc, self.w, self.params["epsilon"], with_fix_sec)
        if self.world > 1:
            dist.all_reduce(cnt, group=self.group)
        return dict(zip(_eng.CLS_NAMES, (int(c) for c in cnt.cpu().numpy())))
