"""srnn_engine -- ctypes binding of libsrnn_b200.so plus the device-resident population arena.

This is the host side of the drop-in: the facade classes in network.py / soup.py / experiment.py (same
names and call signatures as the reference's code/network.py, code/soup.py, code/experiment.py) hold
*views* (population, row) into a `Population`, and every operation on them becomes one call through the
C ABI declared in include/srnn_b200.h.  PyTorch is used for plumbing only: device memory, streams and
torch.distributed -- no torch op computes anything on the hot path.

There is NO CPU fallback: importing works anywhere (so host logic can be tested), but creating a
Population or calling a kernel without the built library and a CUDA device raises immediately.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SRNN_LIB", os.path.join(_HERE, "libsrnn_b200.so"))   # SRNN_LIB: A/B-test another build

KIND_WW, KIND_AGG, KIND_RNN = 0, 1, 2
KIND_NAMES = {KIND_WW: "WeightwiseNeuralNetwork", KIND_AGG: "AggregatingNeuralNetwork", KIND_RNN: "RecurrentNeuralNetwork"}
CLS_NAMES = ("divergent", "fix_zero", "fix_other", "fix_sec", "other")
PERM_NONE, PERM_EXPLICIT, PERM_PHILOX = 0, 1, 2
RUN_FIXED, RUN_UNTIL = 0, 1
ACT_NAMES = (None, "attacking", "learn_from", "train_self", "divergent_dead", "zweo_dead")
FLAG_DIVERGED, FLAG_ZERO, FLAG_FIX1, FLAG_FIX2 = 1, 2, 4, 8

# every symbol include/srnn_b200.h declares (tests check the library exports all of them)
EXPORTS = ("srnn_version", "srnn_last_error_string", "srnn_set_device", "srnn_desc_init", "srnn_shape_supported",
           "srnn_apply", "srnn_forward", "srnn_self_apply", "srnn_classify", "srnn_train", "srnn_init", "srnn_kfv",
           "srnn_soup_ordered", "srnn_soup_exact", "srnn_soup_exact_run", "srnn_soup_sync", "srnn_soup_sync_peers", "srnn_soup_chain_peers", "srnn_soup_push_lists", "srnn_soup_sync_run", "srnn_soup_log_watched", "srnn_assign_uids", "srnn_assign_uids_static", "srnn_fp32_probe",
           "srnn_comm_unique_id", "srnn_comm_init", "srnn_comm_allgather", "srnn_comm_destroy")
UIDS_STATIC_MAX = 1 << 24


class NetDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("width", C.c_int32), ("depth", C.c_int32),
                ("aggregates", C.c_int32), ("n_weights", C.c_int32)]

    def key(self):
        return (self.kind, self.width, self.depth, self.aggregates if self.kind == KIND_AGG else 0)


class SoupParams(C.Structure):
    _fields_ = [("attacking_rate", C.c_double), ("learn_from_rate", C.c_double), ("train", C.c_int32),
                ("learn_from_severity", C.c_int32), ("remove_divergent", C.c_int32), ("remove_zero", C.c_int32),
                ("epsilon", C.c_double), ("lr", C.c_float), ("perm_mode", C.c_int32),
                ("engine_path", C.c_int32), ("dying_cap", C.c_int32), ("dying_w", C.c_void_p),
                ("dying_idx", C.c_void_p), ("dying_count", C.c_void_p)]


class SrnnError(RuntimeError):
    pass


_lib = None


def lib():
    """Load libsrnn_b200.so (built by __graft_entry__.build()).  Fails loudly when it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SrnnError("libsrnn_b200.so not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                            "(there is no CPU fallback)")
        _lib = C.CDLL(LIB_PATH)
        _lib.srnn_last_error_string.restype = C.c_char_p
        _lib.srnn_version.restype = C.c_int
    return _lib


def _chk(rc):
    if rc:
        raise SrnnError("libsrnn_b200 rc=%d: %s" % (rc, lib().srnn_last_error_string().decode()))


def make_desc(kind, width=2, depth=2, aggregates=0):
    d = NetDesc(kind, width, depth, aggregates if kind == KIND_AGG else 0, 0)
    _chk(lib().srnn_desc_init(C.byref(d)))
    return d


def shape_supported(d):
    """0 = invalid shape, 1 = instantiated fast shape (every operation), 2 = generic run-time-shape path (the per-net
    operations: init / apply / forward / self_apply / classify / train; no soups, no KFV sweep)."""
    return int(lib().srnn_shape_supported(C.byref(d)))


def layer_shapes(d):
    """Keras kernel shapes in get_weights() order (code/network.py:226-230, :329-333, :531-535)."""
    io = 4 if d.kind == KIND_WW else (d.aggregates if d.kind == KIND_AGG else 1)
    shapes = []
    for l in range(d.depth + 1):
        fan_in = io if l == 0 else d.width
        fan_out = (1 if d.kind == KIND_WW else io) if l == d.depth else d.width
        shapes.append((fan_in, fan_out))
        if d.kind == KIND_RNN:
            shapes.append((fan_out, fan_out))
    return shapes


def make_soup_params(attacking_rate=0.1, learn_from_rate=0.1, train=0, learn_from_severity=1, remove_divergent=False,
                     remove_zero=False, epsilon=1e-14, lr=0.01, perm_mode=PERM_NONE, engine_path=0, dying=None):
    """dying: optional (records float tensor [cap][n_weights], indices int64 tensor [cap], count int32 tensor [1]) --
    the side channel that keeps the pre-respawn record of every particle that dies (trajectory logging)."""
    sp = SoupParams(float(attacking_rate), float(learn_from_rate), int(train), int(learn_from_severity),
                    int(bool(remove_divergent)), int(bool(remove_zero)), float(epsilon), float(lr), int(perm_mode),
                    int(engine_path), 0, None, None, None)
    if dying is not None:
        rec, idx, cnt = dying
        sp.dying_cap, sp.dying_w, sp.dying_idx, sp.dying_count = int(idx.numel()), rec.data_ptr(), idx.data_ptr(), cnt.data_ptr()
    return sp


def soup_scratch_len(n_total, n):
    """int32 words of scratch srnn_soup_sync / srnn_soup_sync_peers need (include/srnn_b200.h)."""
    return int(n_total) + 4 * int(n) + 16


def _torch():
    import torch
    if not torch.cuda.is_available():
        raise SrnnError("libsrnn_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    return torch


def _ptr(t, ctype=C.c_void_p):
    return C.c_void_p(0) if t is None else C.c_void_p(t.data_ptr())


def _stream():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _sync_device(dev):
    _chk(lib().srnn_set_device(C.c_int(dev.index if dev.index is not None else 0)))


def _idx(t, dev):
    """int32 device index vector (or None)."""
    if t is None:
        return None
    torch = _torch()
    if not isinstance(t, torch.Tensor):
        t = torch.as_tensor(np.ascontiguousarray(t, dtype=np.int32))
    return t.to(device=dev, dtype=torch.int32).contiguous()


class Population:
    """A device arena of `n` net records of one shape: self.w is a float32 CUDA tensor of shape (n, ld)."""

    def __init__(self, desc, n, device=None, ld=None, w=None):
        torch = _torch()
        if not shape_supported(desc):
            raise SrnnError("net shape %r is not valid for libsrnn_b200" % (desc.key(),))
        self.desc = desc
        self.n = int(n)
        self.nw = int(desc.n_weights)
        self.ld = int(ld or self.nw)
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if w is None:
            self.w = torch.zeros((self.n, self.ld), dtype=torch.float32, device=self.device)
        else:
            assert w.dtype == torch.float32 and w.is_cuda and w.is_contiguous() and tuple(w.shape) == (self.n, self.ld)
            self.w = w
        _sync_device(self.device)

    # ---- host <-> device
    @classmethod
    def from_numpy(cls, desc, arr, device=None):
        torch = _torch()
        arr = np.ascontiguousarray(arr, dtype=np.float32)
        pop = cls(desc, arr.shape[0], device=device, ld=arr.shape[1])
        pop.w.copy_(torch.from_numpy(arr))
        return pop

    def numpy(self):
        return self.w[:, :self.nw].cpu().numpy()

    def _args(self):
        return C.byref(self.desc), _ptr(self.w), C.c_int64(self.ld)

    # ---- K7
    def init(self, seed, slot0=0, generation=0, idx=None):
        """Soup.generate_particle for every record (or the records in idx): glorot / orthogonal Keras init."""
        idx = _idx(idx, self.device)
        n = self.n if idx is None else idx.numel()
        _chk(lib().srnn_init(*self._args(), C.c_int64(n), _ptr(idx), C.c_uint64(seed), C.c_uint64(slot0),
                             C.c_uint32(generation), _stream()))
        return self

    # ---- K1
    def self_apply(self, n_steps, mode=RUN_FIXED, eps=1e-14, with_fix_sec=True, want_steps=False, want_cls=False,
                   want_counters=False, want_traj=False):
        """self_attack(n_steps) for every net; mode=RUN_UNTIL adds run_net's stop rule.  Returns a dict of the
        requested outputs (device tensors)."""
        torch = _torch()
        out = {}
        steps = torch.zeros(self.n, dtype=torch.int32, device=self.device) if want_steps else None
        cls = torch.zeros(self.n, dtype=torch.uint8, device=self.device) if want_cls else None
        cnt = torch.zeros(5, dtype=torch.int64, device=self.device) if want_counters else None
        traj = (torch.full((n_steps, self.n, self.ld), float("nan"), dtype=torch.float32, device=self.device)
                if want_traj else None)
        _chk(lib().srnn_self_apply(*self._args(), C.c_int64(self.n), C.c_int32(n_steps), C.c_int32(mode),
                                   C.c_double(eps), C.c_int32(int(with_fix_sec)), _ptr(steps), _ptr(cls), _ptr(cnt),
                                   _ptr(traj), _stream()))
        if want_steps:
            out["steps"] = steps
        if want_cls:
            out["cls"] = cls
        if want_counters:
            out["counters"] = cnt
        if want_traj:
            out["traj"] = traj
        return out

    # ---- K6
    def classify(self, eps, with_fix_sec=True, want_flags=False):
        torch = _torch()
        cls = torch.zeros(self.n, dtype=torch.uint8, device=self.device)
        flags = torch.zeros(self.n, dtype=torch.uint8, device=self.device) if want_flags else None
        cnt = torch.zeros(5, dtype=torch.int64, device=self.device)
        _chk(lib().srnn_classify(*self._args(), C.c_int64(self.n), C.c_double(eps), C.c_int32(int(with_fix_sec)),
                                 _ptr(cls), _ptr(flags), _ptr(cnt), _stream()))
        return (cls, cnt, flags) if want_flags else (cls, cnt)

    # ---- K5
    def train(self, n_epochs=1, lr=0.01, perm_mode=PERM_NONE, perm=None, seed=0, slot0=0, epoch0=0, teacher=None,
              teacher_idx=None, want_loss=True):
        """train() (teacher=None) or learn_from(teacher Population[, teacher_idx]) for every net."""
        torch = _torch()
        loss = torch.full((self.n,), float("nan"), dtype=torch.float64, device=self.device) if want_loss else None
        if perm is not None and not isinstance(perm, torch.Tensor):
            perm = torch.as_tensor(np.ascontiguousarray(perm, dtype=np.uint8))
        if perm is not None:
            perm = perm.to(device=self.device, dtype=torch.uint8).contiguous()
        tidx = _idx(teacher_idx, self.device)
        _chk(lib().srnn_train(*self._args(), C.c_int64(self.n), C.c_int32(n_epochs), C.c_float(lr),
                              C.c_int32(perm_mode), _ptr(perm), C.c_uint64(seed), C.c_uint64(slot0),
                              C.c_uint64(epoch0), _ptr(teacher.w if teacher is not None else None),
                              C.c_int64(teacher.ld if teacher is not None else 0), _ptr(tidx), _ptr(loss), _stream()))
        return loss

    # ---- K2/K3/K4
    def apply_pairs(self, op_idx, in_pop, in_idx, out_pop, out_idx, n=None):
        """out_pop[out_idx[p]] <- f_{self[op_idx[p]]}(in_pop[in_idx[p]])."""
        op_idx, in_idx, out_idx = _idx(op_idx, self.device), _idx(in_idx, self.device), _idx(out_idx, self.device)
        if n is None:
            n = next((t.numel() for t in (op_idx, in_idx, out_idx) if t is not None), self.n)
        _chk(lib().srnn_apply(C.byref(self.desc), _ptr(self.w), C.c_int64(self.ld), _ptr(op_idx), _ptr(in_pop.w),
                              C.c_int64(in_pop.ld), _ptr(in_idx), _ptr(out_pop.w), C.c_int64(out_pop.ld),
                              _ptr(out_idx), C.c_int64(n), _stream()))

    def attack(self, attacker_idx, victim_idx):
        """NeuralNetwork.attack for a list of (attacker, victim) pairs with pairwise distinct records."""
        self.apply_pairs(attacker_idx, self, victim_idx, self, victim_idx)

    # ---- K8
    def forward(self, x, idx=None):
        """NeuralNetwork.apply for a batch of input rows x (n, io) [RNN: (n, n_weights)]; row p uses record idx[p]."""
        torch = _torch()
        if not isinstance(x, torch.Tensor):
            x = torch.as_tensor(np.ascontiguousarray(x, dtype=np.float32))
        x = x.to(device=self.device, dtype=torch.float32).contiguous()
        n = x.shape[0]
        cols = self.nw if self.desc.kind == KIND_RNN else (1 if self.desc.kind == KIND_WW else self.desc.aggregates)
        y = torch.zeros((n, cols), dtype=torch.float32, device=self.device)
        idx = _idx(idx, self.device)
        _chk(lib().srnn_forward(*self._args(), _ptr(idx), _ptr(x), _ptr(y), C.c_int64(n), _stream()))
        return y

    def soup_ordered(self, params, att, lrn, seed, gen, want_logged=False, spares=None):
        torch = _torch()
        att = torch.as_tensor(np.ascontiguousarray(att, dtype=np.int64)).to(self.device)
        lrn = torch.as_tensor(np.ascontiguousarray(lrn, dtype=np.int64)).to(self.device)
        action = torch.zeros(self.n, dtype=torch.uint8, device=self.device)
        loss = torch.full((self.n,), float("nan"), dtype=torch.float64, device=self.device)
        logged = torch.zeros_like(self.w) if want_logged else None
        used = torch.zeros(1, dtype=torch.int32, device=self.device) if spares is not None else None
        _chk(lib().srnn_soup_ordered(C.byref(self.desc), C.byref(params), _ptr(self.w), C.c_int64(self.ld),
                                     C.c_int64(self.n), _ptr(att), _ptr(lrn), C.c_uint64(seed), C.c_uint64(gen),
                                     _ptr(action), _ptr(loss), _ptr(logged), _ptr(spares),
                                     C.c_int64(spares.shape[1] if spares is not None else 0),
                                     C.c_int32(spares.shape[0] if spares is not None else 0), _ptr(used), _stream()))
        out = (action, loss, logged) if want_logged else (action, loss)
        return out + (used,) if spares is not None else out


def soup_sync(desc, params, snap, n_total, w_out, slot0, n, seed, gen, scratch, action=None, loss=None, dead=None):
    """One snapshot-synchronous generation for particles [slot0, slot0+n) (see include/srnn_b200.h)."""
    _chk(lib().srnn_soup_sync(C.byref(desc), C.byref(params), _ptr(snap), C.c_int64(snap.shape[1]), C.c_int64(n_total),
                              _ptr(w_out), C.c_int64(w_out.shape[1]), C.c_int64(slot0), C.c_int64(n), C.c_uint64(seed),
                              C.c_uint64(gen), _ptr(scratch), _ptr(action), _ptr(loss), _ptr(dead), _stream()))


def soup_exact(desc, params, w, seed, gen, scratch, att=None, lrn=None, action=None, loss=None, dead=None,
               want_rounds=True):
    """One generation with the reference's exact sequential semantics, run as a parallel wavefront (in place on w).
    att / lrn: int64 device tensors of targets (or None for Philox draws).  Returns the dependency depth (this waits
    for the stream) unless want_rounds is False (fully asynchronous)."""
    rounds = C.c_int32(0)
    _chk(lib().srnn_soup_exact(C.byref(desc), C.byref(params), _ptr(w), C.c_int64(w.shape[1]), C.c_int64(w.shape[0]),
                               _ptr(att), _ptr(lrn), C.c_uint64(seed), C.c_uint64(gen), _ptr(scratch), _ptr(action),
                               _ptr(loss), _ptr(dead), C.byref(rounds) if want_rounds else None, _stream()))
    return rounds.value


def soup_exact_scratch_len(n):
    """int32 words of scratch srnn_soup_exact / srnn_soup_exact_run need (SRNN_SOUP_EXACT_SCRATCH in the header)."""
    return 11 * int(n) + (int(n) + 1) // 2 + 64


def soup_exact_run(desc, params, w, seed, gen0, n_gens, scratch, action=None, dead=None, rounds=None):
    """n_gens exact-order generations with Philox pairings in ONE asynchronous cooperative launch (in place on w).
    rounds: optional int32 device tensor [n_gens] receiving each generation's dependency depth."""
    _chk(lib().srnn_soup_exact_run(C.byref(desc), C.byref(params), _ptr(w), C.c_int64(w.shape[1]), C.c_int64(w.shape[0]),
                                   C.c_uint64(seed), C.c_uint64(gen0), C.c_int32(n_gens), _ptr(scratch), _ptr(action),
                                   _ptr(dead), _ptr(rounds), _stream()))


def soup_sync_peers(desc, params, shard_ptrs, shard_len, my_shard, ld, w_out, seed, gen, scratch, action=None, loss=None,
                    dead=None, stream=None):
    """One generation of a sharded soup reading partner records straight from the peers' arenas (shard_ptrs = device
    pointers, as ints or a ready ctypes array of void pointers, of every rank's start-of-generation arena mapped into
    this process).  stream: a torch.cuda.Stream to launch on (default: the current stream)."""
    arr = shard_ptrs if isinstance(shard_ptrs, C.Array) else shard_ptr_array(shard_ptrs)
    _chk(lib().srnn_soup_sync_peers(C.byref(desc), C.byref(params), arr, C.c_int32(len(arr)), C.c_int64(shard_len),
                                    C.c_int32(my_shard), C.c_int64(ld), _ptr(w_out), C.c_int64(w_out.shape[1]),
                                    C.c_uint64(seed), C.c_uint64(gen), _ptr(scratch), _ptr(action), _ptr(loss), _ptr(dead),
                                    _stream() if stream is None else C.c_void_p(stream.cuda_stream)))


def soup_log_watched(desc, params, w, n_total, slot0, watch, seed, gen, action, loss, uid_local, ring_w, ring_dying,
                     ring_action, ring_uid, ring_counterpart, ring_loss):
    """One row of the trajectory ring for the watched local indices (include/srnn_b200.h: srnn_soup_log_watched)."""
    _chk(lib().srnn_soup_log_watched(C.byref(desc), C.byref(params), _ptr(w), C.c_int64(w.shape[1]), C.c_int64(n_total),
                                     C.c_int64(slot0), _ptr(watch), C.c_int32(watch.numel()), C.c_uint64(seed),
                                     C.c_uint64(gen), _ptr(action), _ptr(loss), _ptr(uid_local), _ptr(ring_w),
                                     _ptr(ring_dying), _ptr(ring_action), _ptr(ring_uid), _ptr(ring_counterpart),
                                     _ptr(ring_loss), _stream()))


def soup_push_lists(params, n_shards, shard_len, my_shard, seed, gen, heads, nexts, stream=None):
    """First generation of a chained sharded run: this rank's particles push their attacks onto the owners' lists
    (heads / nexts: ctypes arrays of every rank's list pointers, see include/srnn_b200.h)."""
    _chk(lib().srnn_soup_push_lists(C.byref(params), C.c_int32(n_shards), C.c_int64(shard_len), C.c_int32(my_shard),
                                    C.c_uint64(seed), C.c_uint64(gen), heads, nexts,
                                    _stream() if stream is None else C.c_void_p(stream.cuda_stream)))


def soup_chain_peers(desc, params, shard_ptrs, shard_len, my_shard, ld, w_out, seed, gen, heads_cur, nexts_cur, heads_nxt,
                     nexts_nxt, action=None, loss=None, dead=None, stream=None, flags=None, wait_gen=0, scratch=None):
    """One generation of a sharded soup whose attack lists are built by the attackers' ranks: ONE launch, which also
    pushes generation gen+1's attacks (heads_nxt / nexts_nxt, None to skip).  flags / wait_gen: the fused cross-GPU
    signal that replaces the per-generation barrier (include/srnn_b200.h)."""
    arr = shard_ptrs if isinstance(shard_ptrs, C.Array) else shard_ptr_array(shard_ptrs)
    _chk(lib().srnn_soup_chain_peers(C.byref(desc), C.byref(params), arr, C.c_int32(len(arr)), C.c_int64(shard_len),
                                     C.c_int32(my_shard), C.c_int64(ld), _ptr(w_out), C.c_int64(w_out.shape[1]),
                                     C.c_uint64(seed), C.c_uint64(gen), heads_cur, nexts_cur, heads_nxt, nexts_nxt,
                                     flags, C.c_int64(wait_gen), _ptr(scratch), _ptr(action), _ptr(loss), _ptr(dead),
                                     _stream() if stream is None else C.c_void_p(stream.cuda_stream)))


def bind_host_thread_to_gpu(device_index=None):
    """Pin the calling thread (and the pinned host buffers it allocates afterwards: first touch) to the CPUs of the
    NUMA node the GPU hangs off.  One process per GPU on a multi-socket host otherwise stages its host->device copies
    through the other socket.  Returns the CPU set now in force, or None when NVML cannot tell (nothing is changed)."""
    torch = _torch()
    try:
        import pynvml
        idx = torch.cuda.current_device() if device_index is None else int(device_index)
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByUUID("GPU-" + str(torch.cuda.get_device_properties(idx).uuid))
        pynvml.nvmlDeviceSetCpuAffinity(h)
        return sorted(os.sched_getaffinity(0))
    except Exception:
        return None


def shard_ptr_array(shard_ptrs):
    return (C.c_void_p * len(shard_ptrs))(*[C.c_void_p(int(q)) for q in shard_ptrs])


def soup_sync_run(desc, params, w_a, w_b, seed, gen0, n_gens, scratch, action=None, dead=None):
    """n_gens generations of an unsharded soup in ONE call (ping-pong between w_a and w_b; newest = w_b if n_gens odd)."""
    _chk(lib().srnn_soup_sync_run(C.byref(desc), C.byref(params), _ptr(w_a), _ptr(w_b), C.c_int64(w_a.shape[1]),
                                  C.c_int64(w_a.shape[0]), C.c_uint64(seed), C.c_uint64(gen0), C.c_int32(n_gens),
                                  _ptr(scratch), _ptr(action), _ptr(dead), _stream()))


def assign_uids(dead, uid, uid_base, block_tmp, n_dead_out=None):
    _chk(lib().srnn_assign_uids(_ptr(dead), C.c_int64(dead.numel()), C.c_int64(uid_base), _ptr(uid), _ptr(block_tmp),
                                _ptr(n_dead_out), _stream()))


def assign_uids_static(dead, uid, uid_base, n_dead_out=None, n=None):
    """srnn_assign_uids without caller scratch: for hosts that cannot allocate device memory (they pass through
    pointers owned by someone else, or their memory was fixed at start-up).  At most UIDS_STATIC_MAX particles;
    calls on one GPU must not overlap (they share one array inside the library)."""
    _chk(lib().srnn_assign_uids_static(_ptr(dead), C.c_int64(dead.numel() if n is None else n), C.c_int64(uid_base),
                                       _ptr(uid), _ptr(n_dead_out), _stream()))


def kfv(desc, base, n, scale, level, max_steps, eps, seed, slot0=0, want_times=True, want_hist=True,
        want_weights=False, device=None):
    """known-fixpoint-variation trial loop for n perturbed copies of `base` at one noise level."""
    torch = _torch()
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    _sync_device(dev)
    base = np.ascontiguousarray(base, dtype=np.float32)
    assert base.size == desc.n_weights
    ys = torch.zeros(n, dtype=torch.int32, device=dev) if want_times else None
    zs = torch.zeros(n, dtype=torch.int32, device=dev) if want_times else None
    hist = torch.zeros((2, max_steps + 1), dtype=torch.int64, device=dev) if want_hist else None
    wout = torch.zeros((n, desc.n_weights), dtype=torch.float32, device=dev) if want_weights else None
    _chk(lib().srnn_kfv(C.byref(desc), base.ctypes.data_as(C.c_void_p), C.c_int64(n), C.c_double(scale),
                        C.c_uint32(level), C.c_int32(max_steps), C.c_double(eps), C.c_uint64(seed), C.c_uint64(slot0),
                        _ptr(ys), _ptr(zs), _ptr(hist), _ptr(wout), C.c_int64(desc.n_weights), _stream()))
    return dict(ys=ys, zs=zs, hist=hist, w=wout)


class Comm:
    """The C ABI's communicator (srnn_comm_*): NCCL all-gather of record arenas for hosts that do not use
    torch.distributed.  Rank 0 calls Comm.unique_id() and ships the 128 bytes to the other ranks by any means; every rank
    then builds Comm(id, rank, world) with its GPU current (collective)."""

    @staticmethod
    def unique_id():
        buf = (C.c_uint8 * 128)()
        _chk(lib().srnn_comm_unique_id(buf))
        return bytes(buf)

    def __init__(self, uid, rank, world):
        buf = (C.c_uint8 * 128).from_buffer_copy(uid)
        self._h = C.c_void_p()
        _chk(lib().srnn_comm_init(buf, C.c_int32(rank), C.c_int32(world), C.byref(self._h)))
        self.rank, self.world = rank, world

    def allgather(self, send, recv):
        """recv[r * send.numel() ...] = rank r's send (float32 device tensors), asynchronous on the current stream."""
        assert recv.numel() == self.world * send.numel()
        _chk(lib().srnn_comm_allgather(self._h, _ptr(send), _ptr(recv), C.c_int64(send.numel()), _stream()))

    def destroy(self):
        if self._h:
            _chk(lib().srnn_comm_destroy(self._h))
            self._h = C.c_void_p()


def fp32_probe(iters=4096, variant=0, blocks_per_sm=8, device=None):
    """Measured FP32 issue-rate ceiling (lane-FFMA/s) of the current GPU; the roofline denominator of K1/K5."""
    torch = _torch()
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    _sync_device(dev)
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    grid, block = sms * blocks_per_sm, 256
    out = torch.zeros(grid * block, dtype=torch.float32, device=dev)
    it = iters if variant == 0 else (-iters if variant == 1 else -iters - (1 << 24))   # 2 = packed f32x2
    _chk(lib().srnn_fp32_probe(_ptr(out), C.c_int32(grid), C.c_int32(block), C.c_int32(it), _stream()))  # warm-up
    best = 0.0
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _chk(lib().srnn_fp32_probe(_ptr(out), C.c_int32(grid), C.c_int32(block), C.c_int32(it), _stream()))
        e1.record()
        e1.synchronize()
        ms = e0.elapsed_time(e1)
        best = max(best, grid * block * iters * 64.0 / (ms * 1e-3))
    return best
