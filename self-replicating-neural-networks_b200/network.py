"""network -- drop-in for the reference's code/network.py on top of the B200 engine.

Same class names, constructor signatures, method names, return conventions and error behaviour as the
reference (NeuralNetwork, WeightwiseNeuralNetwork, AggregatingNeuralNetwork, RecurrentNeuralNetwork,
ParticleDecorator, TrainingNeuralNetworkDecorator), but a net is not a Keras model: it is ONE RECORD of a
device-resident population arena (srnn_engine.Population), and every operation is a call through the C ABI
of libsrnn_b200.so:

    apply_to_weights / attack / fuck / self_attack   -> srnn_apply, srnn_self_apply
    is_diverged / is_zero / is_fixpoint              -> srnn_classify (predicate flags)
    train / learn_from                               -> srnn_train
    apply (model.predict on one row)                 -> srnn_forward
    fresh Keras initialisation                       -> srnn_init

Per-object calls launch tiny kernels (correct, not fast); populations go through Soup / srnn_engine directly.
Like the reference, every net is linear and bias-free whatever `with_keras_params` says: the reference builds
its Keras layers in __init__ from the default dict and only mutates the dict afterwards
(code/network.py:80, :96-98, :227-230), so the dict is kept for log names only.

Star-import surface kept for the setups: np, copy, os, time, dill, tqdm, K (code/network.py:1-9).
"""
import itertools
import os
import weakref

import numpy as np

from util import *          # noqa: F401,F403  (PrintingObject)
from experiment import *    # noqa: F401,F403  (os, time, dill, tqdm, copy, Experiment ...)
from keras import backend as K  # noqa: F401    (K.clear_session())
import srnn_engine as _eng

os.environ.setdefault("TF_CPP_MIN_LOG_LEVEL", "2")   # the reference sets it; harmless here


# ----------------------------------------------------------------------------------------------- engine state
class _EngineState:
    """Process-wide: the seed of the deterministic Keras-initialiser stand-in, a creation counter (= Philox
    slot of the next fresh net) and one row heap per net shape."""
    seed = int.from_bytes(os.urandom(8), "little") if "SRNN_SEED" not in os.environ else int(os.environ["SRNN_SEED"])
    created = itertools.count()
    heaps = {}
    CHUNK = 1024


def set_seed(seed):
    """Make fresh-net initialisation reproducible (the reference's Keras init is unseeded)."""
    _EngineState.seed = int(seed)
    _EngineState.created = itertools.count()


class _RowHeap:
    """Rows of a growing list of 1024-record device chunks, recycled when their facade object dies."""

    def __init__(self, desc):
        self.desc, self.chunks, self.free = desc, [], []

    def take(self):
        if not self.free:
            chunk = _eng.Population(self.desc, _EngineState.CHUNK)
            self.chunks.append(chunk)
            self.free.extend((chunk, r) for r in reversed(range(_EngineState.CHUNK)))
        chunk, row = self.free.pop()
        view = _eng.Population(self.desc, 1, device=chunk.device, ld=chunk.ld, w=chunk.w[row:row + 1])
        return view, (chunk, row)

    def give_back(self, token):
        self.free.append(token)


def _heap(desc):
    key = desc.key()
    if key not in _EngineState.heaps:
        _EngineState.heaps[key] = _RowHeap(desc)
    return _EngineState.heaps[key]


def _base(net):
    """Strip ParticleDecorator / TrainingNeuralNetworkDecorator wrappers (any stacking order)."""
    seen = 0
    while not isinstance(net, NeuralNetwork) and hasattr(net, "net") and seen < 8:
        net, seen = net.net, seen + 1
    return net if isinstance(net, NeuralNetwork) else None


def _split(flat, shapes):
    out, pos = [], 0
    for r, c in shapes:
        out.append(np.array(flat[pos:pos + r * c], dtype=np.float32).reshape(r, c))
        pos += r * c
    return out


def _flatten(weights):
    return np.concatenate([np.asarray(m, dtype=np.float32).reshape(-1) for m in weights])


class _ModelView:
    """What get_model() returns: the slice of the Keras Sequential API user code touches on a built model."""

    def __init__(self, net):
        self._ref = weakref.ref(net)     # no reference cycle: a dead net frees its arena row by refcount alone

    @property
    def _net(self):
        return self._ref()

    def get_weights(self):
        return self._net.get_weights()

    def set_weights(self, weights):
        return self._net.set_weights(weights)

    def predict(self, x):
        x = np.asarray(x, dtype=np.float32)
        if x.ndim == 3:                       # recurrent input (1, T, 1)
            return self._net._forward(x.reshape(x.shape[0], -1))[..., None]
        return self._net._forward(x)

    def compile(self, **kwargs):
        _check_compile(kwargs)


def _check_compile(params):
    loss, opt = params.get("loss", "mse"), params.get("optimizer", "sgd")
    if loss not in ("mse", "mean_squared_error") or opt != "sgd":
        raise NotImplementedError("the engine implements the reference's compile(loss='mse', optimizer='sgd') only, "
                                  "got loss=%r optimizer=%r" % (loss, opt))


# ----------------------------------------------------------------------------------------------- base class
class NeuralNetwork(PrintingObject):  # noqa: F405
    """A linear, bias-free net whose weights are one record of a device arena (code/network.py:29-163)."""

    KIND = None

    # ---- static helpers on nested weight lists (host bookkeeping, same contracts as the reference)
    @staticmethod
    def weights_to_string(weights):
        lines = []
        for layer in weights:
            lines.append("".join("[ " + "".join(str(w) + " " for w in cell) + "]" for cell in layer))
        return "".join(line + "\n" for line in lines)

    @staticmethod
    def are_weights_diverged(network_weights):
        return any(not np.isfinite(w) for layer in network_weights for cell in layer for w in cell)

    @staticmethod
    def are_weights_within(network_weights, lower_bound, upper_bound):
        return all(lower_bound <= w <= upper_bound for layer in network_weights for cell in layer for w in cell)

    @staticmethod
    def fill_weights(old_weights, new_weights_list):
        it = iter(new_weights_list)
        filled = copy.deepcopy(old_weights)  # noqa: F405
        for layer in filled:
            for cell in layer:
                for j in range(len(cell)):
                    cell[j] = next(it)
        return filled

    # ---- construction
    def __init__(self, **params):
        super().__init__()
        self.params = dict(epsilon=0.00000000000001)
        self.params.update(params)
        self.keras_params = dict(activation='linear', use_bias=False)
        self.states = []
        self._version = 0
        self._flag_cache = None
        self._pending = 0      # self_attack() steps not yet run on the device (see _pop)
        self._row = None

    # ---- the arena row behind this net.  self_attack() of a fused net is DEFERRED: the calls only count up, and the next
    #      access to the row -- any read, write, attack, predicate, training, soup adoption: everything goes through this
    #      property -- runs them as ONE srnn_self_apply launch (n launches of one step and one launch of n steps give the
    #      same bits).  applying-fixpoints.py's `for _ in range(100): net.self_attack()` is one launch instead of 100.
    @property
    def _pop(self):
        if self._pending:
            steps, self._pending = self._pending, 0
            self._row.self_apply(steps)
        return self._row

    @_pop.setter
    def _pop(self, view):
        if self.__dict__.get("_pending") and self.__dict__.get("_row") is not None:
            steps, self._pending = self._pending, 0
            self._row.self_apply(steps)     # the old row keeps the state it was owed
        self._row = view

    def _bind(self, kind, width, depth, aggregates=0):
        self._desc = _eng.make_desc(kind, width, depth, aggregates)
        self._fast_shape = _eng.shape_supported(self._desc) == 1   # else: generic run-time-shape kernels (no soups)
        self._shapes = _eng.layer_shapes(self._desc)
        heap = _heap(self._desc)
        self._pop, token = heap.take()
        self._release = weakref.finalize(self, heap.give_back, token)
        self._pop.init(_EngineState.seed, slot0=next(_EngineState.created))   # fresh Keras-style initialisation
        self.model = _ModelView(self)

    def _rebind(self, view):
        """Move this net into another arena row (a Soup adopts its particles): weights travel, old row is freed."""
        view.w.copy_(self._pop.w)
        self._release()
        self._pop = view
        self._touch()

    def _touch(self):
        self._version += 1

    # ---- dict plumbing
    def get_model(self):
        return self.model

    def get_params(self):
        return self.params

    def get_keras_params(self):
        return self.keras_params

    def with_params(self, **kwargs):
        self.params.update(kwargs)
        return self

    def with_keras_params(self, **kwargs):
        self.keras_params.update(kwargs)     # reported only; the built model stays linear / bias-free
        return self

    # ---- weights in and out (fresh host copies, Keras kernel shapes, float32)
    def get_weights_flat(self):
        return self._pop.numpy()[0].copy()

    def get_weights(self):
        return _split(self.get_weights_flat(), self._shapes)

    def set_weights(self, new_weights):
        flat = _flatten(new_weights)
        if flat.size != self._desc.n_weights:
            raise ValueError("expected %d weights, got %d" % (self._desc.n_weights, flat.size))
        import torch
        self._pop.w[0, :flat.size].copy_(torch.from_numpy(flat))
        self._touch()

    # ---- the operator
    def _forward(self, rows):
        return self._pop.forward(np.ascontiguousarray(rows, dtype=np.float32),
                                 idx=np.zeros(len(rows), np.int32)).cpu().numpy()

    def _apply_flat(self, flat):
        """f_self(flat) for a same-shaped weight vector, on the device."""
        if flat.size != self._desc.n_weights:
            raise NotImplementedError("cross-shape application is undefined for this net kind in the engine")
        data = _eng.Population.from_numpy(self._desc, flat.reshape(1, -1))
        self._pop.apply_pairs(None, data, None, data, None, n=1)
        return data.numpy()[0]

    def apply_to_weights(self, old_weights):
        shapes = [np.asarray(m).shape for m in old_weights]
        return _split(self._apply_flat(_flatten(old_weights)), shapes)

    def apply_to_network(self, other_network):
        return self.apply_to_weights(other_network.get_weights())

    def _fused(self):
        """True when this net's operator is one of the engine's fused kernels (no user override / hooks)."""
        return type(self).apply_to_weights in _FUSED_OPERATORS

    def attack(self, other_network):
        victim = _base(other_network)
        if self._fused() and victim is not None and victim._desc.key() == self._desc.key():
            self._pop.apply_pairs(None, victim._pop, None, victim._pop, None, n=1)   # in place on the device
            victim._touch()
        else:
            other_network.set_weights(self.apply_to_network(other_network))
        return self

    def fuck(self, other_network):
        self.set_weights(self.apply_to_network(other_network))
        return self

    def self_attack(self, iterations=1):
        if self._fused():
            self._pending += int(iterations)
            self._touch()
        else:
            for _ in range(iterations):
                self.attack(self)
        return self

    def meet(self, other_network):
        return self.attack(copy.deepcopy(other_network))  # noqa: F405  (the clone is discarded: no lasting effect)

    # ---- predicates
    def _eps(self, epsilon):
        return epsilon or self.get_params().get('epsilon')

    def _flags(self, eps):
        key = (self._version, float(eps))
        if self._flag_cache is None or self._flag_cache[0] != key:
            _, _, flags = self._pop.classify(eps, want_flags=True)
            self._flag_cache = (key, int(flags[0].item()))
        return self._flag_cache[1]

    def is_diverged(self):
        return bool(self._flags(self.params.get('epsilon')) & _eng.FLAG_DIVERGED)

    def is_zero(self, epsilon=None):
        return bool(self._flags(self._eps(epsilon)) & _eng.FLAG_ZERO)

    def is_fixpoint(self, degree=1, epsilon=None):
        assert degree >= 1, "degree must be >= 1"
        eps = self._eps(epsilon)
        if degree <= 2 and self._fused():
            return bool(self._flags(eps) & (_eng.FLAG_FIX1 if degree == 1 else _eng.FLAG_FIX2))
        old = self.get_weights()
        new = copy.deepcopy(old)  # noqa: F405
        for _ in range(degree):
            new = self.apply_to_weights(new)
        if NeuralNetwork.are_weights_diverged(new):
            return False
        return all(abs(n - o) < eps for ln, lo in zip(new, old) for cn, co in zip(ln, lo) for n, o in zip(cn, co))

    def _engine_class_name(self):
        """FixpointExperiment.count's decision list in one launch."""
        cls, _ = self._pop.classify(self.params.get('epsilon'))
        return _eng.CLS_NAMES[int(cls[0].item())]

    def _engine_run_until(self, step_limit, log_states=False):
        """FixpointExperiment.run_net's loop in one launch; returns (steps, states after each step or None)."""
        out = self._pop.self_apply(int(step_limit), mode=_eng.RUN_UNTIL, eps=self.params.get('epsilon'),
                                   want_steps=True, want_traj=log_states)
        self._touch()
        steps = int(out["steps"][0].item())
        traj = out["traj"][:steps, 0, :self._desc.n_weights].cpu().numpy() if log_states else None
        return steps, traj

    def repr_weights(self, weights=None):
        return self.weights_to_string(weights or self.get_weights())

    def print_weights(self, weights=None):
        print(self.repr_weights(weights))

    def __deepcopy__(self, memo):
        twin = self.__class__.__new__(self.__class__)
        memo[id(self)] = twin
        for k, v in self.__dict__.items():
            if k in ("_row", "_pending", "_release", "model", "_flag_cache"):
                continue
            twin.__dict__[k] = copy.deepcopy(v, memo)  # noqa: F405
        heap = _heap(self._desc)
        twin._pending, twin._row = 0, None
        twin._pop, token = heap.take()
        twin._release = weakref.finalize(twin, heap.give_back, token)
        twin._pop.w.copy_(self._pop.w)
        twin.model, twin._flag_cache = _ModelView(twin), None
        return twin


# ----------------------------------------------------------------------------------------------- particles
class ParticleDecorator:
    """Identity + trajectory log around a net (code/network.py:166-210): uid from a process-wide counter, a
    `states` list of {class, weights, time, action, counterpart, ...}; non-finite states are skipped."""
    next_uid = 0

    def __init__(self, net):
        self.uid = self.__class__.next_uid
        self.__class__.next_uid += 1
        self.net = net
        self.states = []
        self.save_state(time=0, action='init', counterpart=None)

    def __getattr__(self, name):
        if name == "net" or name.startswith("__"):
            raise AttributeError(name)
        return getattr(self.net, name)

    def get_uid(self):
        return self.uid

    def _state_of(self, weights, **kwargs):
        if not np.isfinite(weights).all():
            return None
        state = {'class': self.net.__class__.__name__, 'weights': weights}
        state.update(kwargs)
        return state

    def make_state(self, **kwargs):
        return self._state_of(self.net.get_weights_flat(), **kwargs)

    def _log(self, state):
        if state is not None:
            self.states += [state]

    def save_state(self, **kwargs):
        self._log(self.make_state(**kwargs))

    def update_state(self, number, **kwargs):
        raise NotImplementedError('Result is vague')

    def get_states(self):
        return self.states

    def _engine_run_until(self, step_limit, log_states=False):
        steps, traj = self.net._engine_run_until(step_limit, log_states)
        if log_states:
            for i in range(steps):
                self._log(self._state_of(traj[i].copy(), time=i + 1))
        return steps, traj


# ----------------------------------------------------------------------------------------------- weightwise
class WeightwiseNeuralNetwork(NeuralNetwork):
    """MLP 4 -> width x depth -> 1 applied to every (weight, layer, row, column) point of a weight list
    (code/network.py:213-289)."""
    KIND = _eng.KIND_WW

    @staticmethod
    def normalize_id(value, norm):
        return float(value) / float(norm) if norm > 1 else float(value)

    def __init__(self, width, depth, **kwargs):
        super().__init__(**kwargs)
        self.width, self.depth = width, depth
        self._bind(_eng.KIND_WW, width, depth)

    def apply(self, *inputs):
        return self._forward(np.asarray(inputs[:4], dtype=np.float32).reshape(1, 4))[0][0]

    @classmethod
    def compute_all_duplex_weight_points(cls, old_weights):
        raw, normal = [], []
        top_layer = len(old_weights) - 1
        for li, layer in enumerate(old_weights):
            top_cell = len(layer) - 1
            for ci, cell in enumerate(layer):
                top_weight = len(cell) - 1
                for wi, weight in enumerate(cell):
                    raw.append([weight, li, ci, wi])
                    normal.append([weight, cls.normalize_id(li, top_layer), cls.normalize_id(ci, top_cell),
                                   cls.normalize_id(wi, top_weight)])
        return raw, normal

    @classmethod
    def compute_all_weight_points(cls, all_weights):
        return cls.compute_all_duplex_weight_points(all_weights)[0]

    @classmethod
    def compute_all_normal_weight_points(cls, all_weights):
        return cls.compute_all_duplex_weight_points(all_weights)[1]

    def apply_to_weights(self, old_weights):
        flat = _flatten(old_weights)
        if flat.size != self._desc.n_weights:
            # the reference writes results into a copy of SELF's weights and raises on a shape mismatch
            raise IndexError("weightwise application needs attacker and victim of the same shape")
        return _split(self._apply_flat(flat), self._shapes)

    def compute_samples(self):
        pts = np.asarray(self.compute_all_normal_weight_points(self.get_weights()))
        return pts, pts[:, 0]


# ----------------------------------------------------------------------------------------------- aggregating
class AggregatingNeuralNetwork(NeuralNetwork):
    """MLP a -> width x depth -> a on `a` block aggregates of the flat weight list (code/network.py:292-439)."""
    KIND = _eng.KIND_AGG

    @staticmethod
    def aggregate_average(weights):
        total, count = 0, 0
        for weight in weights:
            total += float(weight)
            count += 1
        return total / float(count)

    @staticmethod
    def aggregate_max(weights):
        best = weights[0]
        for weight in weights:
            best = weight > best and weight or best
        return best

    @staticmethod
    def deaggregate_identically(aggregate, amount):
        return [aggregate] * amount

    @staticmethod
    def shuffle_not(weights_list):
        return weights_list

    @staticmethod
    def shuffle_random(weights_list):
        import random
        random.shuffle(weights_list)
        return weights_list

    def __init__(self, aggregates, width, depth, **kwargs):
        super().__init__(**kwargs)
        self.aggregates, self.width, self.depth = aggregates, width, depth
        self._bind(_eng.KIND_AGG, width, depth, aggregates)

    def get_aggregator(self):
        return self.params.get('aggregator', self.aggregate_average)

    def get_deaggregator(self):
        return self.params.get('deaggregator', self.deaggregate_identically)

    def get_shuffler(self):
        return self.params.get('shuffler', self.shuffle_not)

    def _default_hooks(self):
        return not any(k in self.params for k in ('aggregator', 'deaggregator', 'shuffler'))

    def _fused(self):
        return super()._fused() and self._default_hooks()

    def get_amount_of_weights(self):
        return int(self._desc.n_weights)

    def apply(self, *inputs):
        return self._forward(np.asarray(inputs[:self.aggregates], dtype=np.float32).reshape(1, -1))[0]

    def apply_to_weights(self, old_weights):
        flat = _flatten(old_weights)
        if self._default_hooks() and flat.size == self._desc.n_weights:
            return _split(self._apply_flat(flat), [np.asarray(m).shape for m in old_weights])   # fused device path
        # user-supplied aggregator / deaggregator / shuffler (params hooks) or a victim of another size:
        # the hooks are Python callables, so they run on the host around the device forward pass
        size = self.get_amount_of_weights() // self.aggregates
        collections, leftovers = self.collect_weights(old_weights, size)
        new_aggs = self.apply(*[self.get_aggregator()(c) for c in collections])
        out = []
        for i, agg in enumerate(new_aggs):
            out += self.get_deaggregator()(agg, size + leftovers if i == self.aggregates - 1 else size)
        return self.fill_weights(old_weights, self.get_shuffler()(out))

    @staticmethod
    def collect_weights(all_weights, collection_size):
        flat = [w for layer in all_weights for cell in layer for w in cell]
        full = len(flat) // collection_size
        collections = [flat[i * collection_size:(i + 1) * collection_size] for i in range(full)]
        rest = flat[full * collection_size:]
        collections[-1] += rest
        return collections, len(rest)

    def get_collected_weights(self):
        return self.collect_weights(self.get_weights(), self.get_amount_of_weights() // self.aggregates)

    def get_aggregated_weights(self):
        collections, leftovers = self.get_collected_weights()
        return [self.get_aggregator()(c) for c in collections], leftovers

    def compute_samples(self):
        aggs, _ = self.get_aggregated_weights()
        sample = np.transpose(np.array([[aggs[i]] for i in range(self.aggregates)]))
        return [sample], [sample]

    def is_fixpoint_after_aggregation(self, degree=1, epsilon=None):
        assert degree >= 1, "degree must be >= 1"
        eps = self._eps(epsilon)
        old_aggs, _ = self.get_aggregated_weights()
        new = copy.deepcopy(self.get_weights())  # noqa: F405
        for _ in range(degree):
            new = self.apply_to_weights(new)
        if NeuralNetwork.are_weights_diverged(new):
            return False
        collections, _ = self.collect_weights(new, self.get_amount_of_weights() // self.aggregates)
        new_aggs = [self.get_aggregator()(c) for c in collections]
        for old, cur in zip(old_aggs, new_aggs):
            if abs(cur - old) >= eps:
                return False, new_aggs
        return True, new_aggs


class FFTNeuralNetwork(NeuralNetwork):
    """Present in the reference (code/network.py:442-521) but broken there ("FFT doesn't work though",
    code/setups/fixpoint-density.py:35) and never run: out of scope."""

    def __init__(self, aggregates, width, depth, **kwargs):
        raise NotImplementedError("FFTNeuralNetwork is out of scope: it does not work in the reference either")


# ----------------------------------------------------------------------------------------------- recurrent
class RecurrentNeuralNetwork(NeuralNetwork):
    """Stacked linear SimpleRNN 1 -> width x depth -> 1 reading the flat weight list as a sequence
    (code/network.py:524-574)."""
    KIND = _eng.KIND_RNN

    def __init__(self, width, depth, **kwargs):
        super().__init__(**kwargs)
        self.features, self.width, self.depth = 1, width, depth
        self._bind(_eng.KIND_RNN, width, depth)

    def apply(self, *inputs):
        if len(inputs) != self._desc.n_weights:
            raise NotImplementedError("the engine's recurrent forward takes sequences as long as the weight list")
        return self._forward(np.asarray(inputs, dtype=np.float32).reshape(1, -1))[0]

    def compute_samples(self):
        sample = self.get_weights_flat()[None, ..., None]
        return sample, sample


_FUSED_OPERATORS = {NeuralNetwork.apply_to_weights, WeightwiseNeuralNetwork.apply_to_weights,
                    AggregatingNeuralNetwork.apply_to_weights}


# ----------------------------------------------------------------------------------------------- training
class TrainingNeuralNetworkDecorator:
    """compile(loss='mse', optimizer='sgd') + train / learn_from (code/network.py:577-626): one Keras epoch of
    SGD(lr=0.01), batch size 1, per call, fused on the device.  Keras shuffles the samples of an epoch with the
    global NumPy RNG; so does this class (np.random.permutation), which makes `np.random.seed` meaningful."""

    def __init__(self, net, **kwargs):
        self.net = net
        self.compile_params = dict(loss='mse', optimizer='sgd')
        self.model_compiled = False

    def __getattr__(self, name):
        if name == "net" or name.startswith("__"):
            raise AttributeError(name)
        return getattr(self.net, name)

    def with_params(self, **kwargs):
        self.net.with_params(**kwargs)
        return self

    def with_keras_params(self, **kwargs):
        self.net.with_keras_params(**kwargs)
        return self

    def get_compile_params(self):
        return self.compile_params

    def with_compile_params(self, **kwargs):
        self.compile_params.update(kwargs)
        return self

    def compile_model(self, **kwargs):
        merged = copy.deepcopy(self.compile_params)  # noqa: F405
        merged.update(kwargs)
        return _check_compile(merged)

    def compiled(self, **kwargs):
        if not self.model_compiled:
            self.compile_model(**kwargs)
            self.model_compiled = True
        return self

    def _fit_one_epoch(self, teacher, batchsize):
        if batchsize != 1:
            raise NotImplementedError("the engine implements the reference's batch_size=1 only")
        me = _base(self)
        ns = me._desc.n_weights if me._desc.kind == _eng.KIND_WW else 1
        perm = np.random.permutation(ns).astype(np.uint8)[None, :] if ns > 1 else None
        loss = me._pop.train(1, lr=0.01, perm_mode=_eng.PERM_EXPLICIT if ns > 1 else _eng.PERM_NONE, perm=perm,
                             teacher=None if teacher is None else teacher._pop)
        me._touch()
        return float(loss[0].item())

    def train(self, batchsize=1, store_states=True, epoch=0):
        self.compiled()
        loss = self._fit_one_epoch(None, batchsize)
        if store_states:
            # SaveStateCallback.on_epoch_end (code/network.py:21-25): Keras reports epoch index = initial_epoch
            self.save_state(time=epoch + epoch, action='train_self', counterpart=None)
        return loss

    def learn_from(self, other_network, batchsize=1):
        self.compiled()
        other_network.compiled()
        teacher = _base(other_network.net)
        me = _base(self)
        if teacher is None or teacher._desc.key() != me._desc.key():
            raise ValueError("learn_from needs a teacher of the same net shape")
        return self._fit_one_epoch(teacher, batchsize)
