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
        if not bases or any(b is None or not b._fused() for b in bases):
            return False
        if len({b._desc.key() for b in bases}) != 1:
            return False
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


class PopulationSoup:
    """Population-scale soup of ONE net shape, device-resident, optionally sharded by particle index over the ranks
    of a torch.distributed process group (one rank per GPU).  One generation =
        [all-gather of the record arena over NVLink when sharded] -> srnn_soup_sync on the local slice
    Pairings, SGD sample orders and respawns are Philox functions of (seed, global particle index, generation),
    so the result does not depend on how the population is sharded."""

    def __init__(self, kind, size, width=2, depth=2, aggregates=0, seed=0, group=None, device=None, **params):
        import torch
        import torch.distributed as dist
        self.desc = _eng.make_desc(kind, width, depth, aggregates)
        self.group = group
        self.world = dist.get_world_size(group) if (group is not None or (dist.is_available() and dist.is_initialized())) else 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        if size % self.world:
            raise ValueError("population size must be a multiple of the number of ranks")
        self.size, self.n_local = int(size), int(size) // self.world
        self.slot0 = self.rank * self.n_local
        self.seed, self.time = int(seed), 0
        self.params = dict(attacking_rate=0.1, learn_from_rate=0.1, train=0, learn_from_severity=1,
                           remove_divergent=False, remove_zero=False, epsilon=1e-4, perm_mode=_eng.PERM_PHILOX)
        self.params.update(params)
        self.local = _eng.Population(self.desc, self.n_local, device=device)
        dev = self.local.device
        nw = self.desc.n_weights
        self._snap = torch.zeros((self.size, nw), dtype=torch.float32, device=dev) if self.world > 1 else None
        self._next = torch.zeros((self.n_local, nw), dtype=torch.float32, device=dev)
        self._scratch = torch.zeros(self.size + self.n_local, dtype=torch.int32, device=dev)
        self.dead = torch.zeros(self.n_local, dtype=torch.uint8, device=dev)
        self.action = torch.zeros(self.n_local, dtype=torch.uint8, device=dev)
        self.uid = torch.arange(self.slot0, self.slot0 + self.n_local, dtype=torch.int64, device=dev)
        self._uid_tmp = torch.zeros((self.n_local + 1023) // 1024 + 1, dtype=torch.int64, device=dev)
        self._n_dead = torch.zeros(1, dtype=torch.int64, device=dev)
        self.next_uid = self.size
        self.track_uids = False

    def with_params(self, **kwargs):
        self.params.update(kwargs)
        return self

    def seed_particles(self):
        """Soup.seed: fresh Keras-style nets for every local particle."""
        self.local.init(self.seed, slot0=self.slot0, generation=0)
        return self

    def _soup_params(self):
        p = self.params
        return _eng.make_soup_params(**{k: p[k] for k in ("attacking_rate", "learn_from_rate", "train",
                                                          "learn_from_severity", "remove_divergent", "remove_zero",
                                                          "epsilon", "perm_mode")})

    def evolve(self, iterations=1):
        import torch
        import torch.distributed as dist
        sp = self._soup_params()
        if self.world == 1 and not self.track_uids and iterations > 1:
            # unsharded: all generations in one C call (no per-generation host overhead)
            _eng.soup_sync_run(self.desc, sp, self.local.w, self._next, self.seed, self.time + 1, iterations,
                               self._scratch, action=self.action, dead=self.dead)
            self.time += iterations
            if iterations & 1:
                self.local.w, self._next = self._next, self.local.w
            return self
        for _ in range(iterations):
            self.time += 1
            if self.world > 1:
                dist.all_gather_into_tensor(self._snap, self.local.w, group=self.group)   # NCCL over NVLink
                snap = self._snap
            else:
                snap = self.local.w
            _eng.soup_sync(self.desc, sp, snap, self.size, self._next, self.slot0, self.n_local, self.seed, self.time,
                           self._scratch, action=self.action, dead=self.dead)
            self.local.w, self._next = self._next, self.local.w
            if self.track_uids:
                self._assign_uids()
        return self

    def _assign_uids(self):
        """Deterministic global uids for respawned particles: rank offsets by an all-gather of the per-rank counts."""
        import torch
        import torch.distributed as dist
        if self.world > 1:
            counts = torch.zeros(self.world, dtype=torch.int64, device=self.local.device)
            mine = self.dead.sum(dtype=torch.int64).reshape(1)
            dist.all_gather_into_tensor(counts, mine, group=self.group)
            c = counts.cpu()
            base, total = self.next_uid + int(c[:self.rank].sum()), int(c.sum())
        else:
            base, total = self.next_uid, None
        _eng.assign_uids(self.dead, self.uid, base, self._uid_tmp, self._n_dead)
        self.next_uid += int(self._n_dead.item()) if total is None else total

    def count(self, with_fix_sec=True):
        """Soup.count over the whole (possibly sharded) population."""
        import torch.distributed as dist
        _, cnt = self.local.classify(self.params["epsilon"], with_fix_sec=with_fix_sec)
        if self.world > 1:
            dist.all_reduce(cnt, group=self.group)
        return dict(zip(_eng.CLS_NAMES, (int(c) for c in cnt.cpu().numpy())))
