"""ctypes binding of the CPU ORACLE (oracle/srnn_oracle.c) -- test infrastructure only.

Importers allowed: tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs.  The product package never imports this module.

All population arguments are numpy float32 arrays of shape (n_nets, ld) -- one record per net, the
same arena layout the CUDA library uses (include/srnn_types.h).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libsrnn_oracle.so")

KIND_WW, KIND_AGG, KIND_RNN = 0, 1, 2
CLS_NAMES = ("divergent", "fix_zero", "fix_other", "fix_sec", "other")
PERM_NONE, PERM_EXPLICIT, PERM_PHILOX = 0, 1, 2
RUN_FIXED, RUN_UNTIL = 0, 1
ACT_NAMES = (None, "attacking", "learn_from", "train_self", "divergent_dead", "zweo_dead")


class NetDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("width", C.c_int32), ("depth", C.c_int32),
                ("aggregates", C.c_int32), ("n_weights", C.c_int32)]


class SoupParams(C.Structure):
    _fields_ = [("attacking_rate", C.c_double), ("learn_from_rate", C.c_double), ("train", C.c_int32),
                ("learn_from_severity", C.c_int32), ("remove_divergent", C.c_int32), ("remove_zero", C.c_int32),
                ("epsilon", C.c_double), ("lr", C.c_float), ("perm_mode", C.c_int32),
                ("engine_path", C.c_int32), ("dying_cap", C.c_int32), ("dying_w", C.c_void_p),
                ("dying_idx", C.c_void_p), ("dying_count", C.c_void_p)]


def build(force=False):
    src = os.path.join(_HERE, "srnn_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
    return _lib


def desc(kind, width=2, depth=2, aggregates=0):
    d = NetDesc(kind, width, depth, aggregates, 0)
    rc = lib().srnn_oracle_desc_init(C.byref(d))
    if rc:
        raise ValueError("srnn_oracle_desc_init rc=%d" % rc)
    return d


def ww(width=2, depth=2):
    return desc(KIND_WW, width, depth)


def agg(aggregates=4, width=2, depth=2):
    return desc(KIND_AGG, width, depth, aggregates)


def rnn(width=2, depth=2):
    return desc(KIND_RNN, width, depth)


def soup_params(attacking_rate=0.1, learn_from_rate=0.1, train=0, learn_from_severity=1, remove_divergent=False,
                remove_zero=False, epsilon=1e-14, lr=0.01, perm_mode=PERM_NONE, engine_path=0):
    return SoupParams(attacking_rate, learn_from_rate, train, learn_from_severity, int(bool(remove_divergent)),
                      int(bool(remove_zero)), epsilon, lr, perm_mode, engine_path, 0, None, None, None)


def _f32(a):
    assert a.dtype == np.float32 and a.ndim == 2 and a.flags.c_contiguous, (a.dtype, a.shape)
    return a.ctypes.data_as(C.POINTER(C.c_float)), C.c_int64(a.shape[1])


def _i32(a):
    if a is None:
        return None
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a, a.ctypes.data_as(C.POINTER(C.c_int32))


def _chk(rc):
    if rc:
        raise RuntimeError("oracle rc=%d" % rc)


def apply(d, op_w, op_idx, in_w, in_idx, out_w, out_idx, n):
    keep = [_i32(op_idx), _i32(in_idx), _i32(out_idx)]
    ptrs = [k[1] if k else None for k in keep]
    (po, lo), (pi, li), (pw, lw) = _f32(op_w), _f32(in_w), _f32(out_w)
    _chk(lib().srnn_oracle_apply(C.byref(d), po, lo, ptrs[0], pi, li, ptrs[1], pw, lw, ptrs[2], C.c_int64(n)))


def self_apply(d, w, n_steps, mode=RUN_FIXED, eps=1e-14, traj=False, n=None):
    n = w.shape[0] if n is None else n
    steps = np.zeros(n, np.int32)
    tr = np.full((n_steps, n, w.shape[1]), np.nan, np.float32) if traj else None
    pw, ld = _f32(w)
    _chk(lib().srnn_oracle_self_apply(C.byref(d), pw, ld, C.c_int64(n), C.c_int32(n_steps), C.c_int32(mode),
                                      C.c_double(eps), steps.ctypes.data_as(C.POINTER(C.c_int32)),
                                      tr.ctypes.data_as(C.POINTER(C.c_float)) if traj else None))
    return (steps, tr) if traj else steps


def self_apply_short_form(d, w, n_steps, exact_last=True, n=None):
    """Test infrastructure: n_steps self-applications with steps 1 .. n-1 in the CUDA kernels' short form (a product
    flushed to -0 keeps its sign in the output sum) and the last one exact (exact_last=False: short form throughout)
    -- tests/test_oracle_zero_signs.py."""
    n = w.shape[0] if n is None else n
    pw, ld = _f32(w)
    _chk(lib().srnn_oracle_self_apply_short_form(C.byref(d), pw, ld, C.c_int64(n), C.c_int32(n_steps),
                                                 C.c_int32(1 if exact_last else 0)))


def classify(d, w, eps, with_fix_sec=True, n=None):
    n = w.shape[0] if n is None else n
    cls = np.zeros(n, np.uint8)
    cnt = np.zeros(5, np.int64)
    pw, ld = _f32(w)
    _chk(lib().srnn_oracle_classify(C.byref(d), pw, ld, C.c_int64(n), C.c_double(eps), C.c_int32(int(with_fix_sec)),
                                    cls.ctypes.data_as(C.POINTER(C.c_uint8)),
                                    cnt.ctypes.data_as(C.POINTER(C.c_int64))))
    return cls, cnt


def predicates(d, w, eps, n=None):
    n = w.shape[0] if n is None else n
    fl = np.zeros(n, np.uint8)
    pw, ld = _f32(w)
    _chk(lib().srnn_oracle_predicates(C.byref(d), pw, ld, C.c_int64(n), C.c_double(eps),
                                      fl.ctypes.data_as(C.POINTER(C.c_uint8))))
    return fl


def philox(c, k):
    out = np.zeros(4, np.uint32)
    lib().srnn_oracle_philox(*[C.c_uint32(int(x)) for x in c], *[C.c_uint32(int(x)) for x in k],
                             out.ctypes.data_as(C.POINTER(C.c_uint32)))
    return out


def init(d, w, seed, slot0=0, generation=0, idx=None, n=None):
    n = (w.shape[0] if idx is None else len(idx)) if n is None else n
    keep = _i32(idx)
    pw, ld = _f32(w)
    _chk(lib().srnn_oracle_init(C.byref(d), pw, ld, C.c_int64(n), keep[1] if keep else None, C.c_uint64(seed),
                                C.c_uint64(slot0), C.c_uint32(generation)))


def perm(seed, slot, epoch, n):
    out = np.zeros(n, np.uint8)
    lib().srnn_oracle_perm(C.c_uint64(seed), C.c_uint64(slot), C.c_uint64(epoch), C.c_int32(n),
                           out.ctypes.data_as(C.POINTER(C.c_uint8)))
    return out


def train(d, w, n_epochs, lr=0.01, perm_mode=PERM_NONE, perm=None, seed=0, slot0=0, epoch0=0, teacher_w=None,
          teacher_idx=None, n=None):
    n = w.shape[0] if n is None else n
    loss = np.full(n, np.nan, np.float64)
    pw, ld = _f32(w)
    if perm is not None:
        perm = np.ascontiguousarray(perm, dtype=np.uint8)
    keep = _i32(teacher_idx)
    pt, lt = _f32(teacher_w) if teacher_w is not None else (None, C.c_int64(0))
    _chk(lib().srnn_oracle_train(C.byref(d), pw, ld, C.c_int64(n), C.c_int32(n_epochs), C.c_float(lr),
                                 C.c_int32(perm_mode), perm.ctypes.data_as(C.POINTER(C.c_uint8)) if perm is not None else None,
                                 C.c_uint64(seed), C.c_uint64(slot0), C.c_uint64(epoch0), pt, lt,
                                 keep[1] if keep else None, loss.ctypes.data_as(C.POINTER(C.c_double))))
    return loss


def kfv(d, base, n, scale, level, max_steps, eps, seed, slot0=0, want_weights=False):
    base = np.ascontiguousarray(base, dtype=np.float32)
    ys = np.zeros(n, np.int32)
    zs = np.zeros(n, np.int32)
    wout = np.zeros((n, d.n_weights), np.float32) if want_weights else None
    _chk(lib().srnn_oracle_kfv(C.byref(d), base.ctypes.data_as(C.POINTER(C.c_float)), C.c_int64(n), C.c_double(scale),
                               C.c_uint32(level), C.c_int32(max_steps), C.c_double(eps), C.c_uint64(seed),
                               C.c_uint64(slot0), ys.ctypes.data_as(C.POINTER(C.c_int32)),
                               zs.ctypes.data_as(C.POINTER(C.c_int32)),
                               wout.ctypes.data_as(C.POINTER(C.c_float)) if want_weights else None, C.c_int64(d.n_weights)))
    return (ys, zs, wout) if want_weights else (ys, zs)


def soup_draw(p, seed, gen, n_total, slot0=0, n=None):
    n = n_total if n is None else n
    att = np.zeros(n, np.int64)
    lrn = np.zeros(n, np.int64)
    lib().srnn_oracle_soup_draw(C.byref(p), C.c_uint64(seed), C.c_uint64(gen), C.c_int64(n_total), C.c_int64(slot0),
                                C.c_int64(n), att.ctypes.data_as(C.POINTER(C.c_int64)),
                                lrn.ctypes.data_as(C.POINTER(C.c_int64)))
    return att, lrn


def forward(d, w, x, idx=None):
    x = np.ascontiguousarray(x, dtype=np.float32)
    n = x.shape[0]
    io_out = x.shape[1] if d.kind == KIND_RNN else (1 if d.kind == KIND_WW else d.aggregates)
    y = np.zeros((n, io_out), np.float32)
    keep = _i32(idx)
    pw, ld = _f32(w)
    _chk(lib().srnn_oracle_forward(C.byref(d), pw, ld, keep[1] if keep else None, x.ctypes.data_as(C.POINTER(C.c_float)),
                                   y.ctypes.data_as(C.POINTER(C.c_float)), C.c_int64(n)))
    return y


def soup_ordered(d, p, w, att, lrn, seed, gen, n=None, want_logged=False, spares=None):
    n = w.shape[0] if n is None else n
    att = np.ascontiguousarray(att, dtype=np.int64)
    lrn = np.ascontiguousarray(lrn, dtype=np.int64)
    action = np.zeros(n, np.uint8)
    loss = np.full(n, np.nan, np.float64)
    logged = np.zeros_like(w) if want_logged else None
    used = C.c_int32(0)
    if spares is not None:
        spares = np.ascontiguousarray(spares, dtype=np.float32)
    pw, ld = _f32(w)
    _chk(lib().srnn_oracle_soup_ordered(C.byref(d), C.byref(p), pw, ld, C.c_int64(n),
                                        att.ctypes.data_as(C.POINTER(C.c_int64)),
                                        lrn.ctypes.data_as(C.POINTER(C.c_int64)), C.c_uint64(seed), C.c_uint64(gen),
                                        action.ctypes.data_as(C.POINTER(C.c_uint8)),
                                        loss.ctypes.data_as(C.POINTER(C.c_double)),
                                        logged.ctypes.data_as(C.POINTER(C.c_float)) if want_logged else None,
                                        spares.ctypes.data_as(C.POINTER(C.c_float)) if spares is not None else None,
                                        C.c_int64(spares.shape[1] if spares is not None else 0),
                                        C.c_int32(spares.shape[0] if spares is not None else 0), C.byref(used)))
    if spares is not None:
        return (action, loss, logged, used.value) if want_logged else (action, loss, used.value)
    return (action, loss, logged) if want_logged else (action, loss)


def soup_sync(d, p, snap, n_total, w_out, slot0, n, seed, gen):
    action = np.zeros(n, np.uint8)
    loss = np.full(n, np.nan, np.float64)
    natt = C.c_int64(0)
    ps, ls = _f32(snap)
    po, lo = _f32(w_out)
    _chk(lib().srnn_oracle_soup_sync(C.byref(d), C.byref(p), ps, ls, C.c_int64(n_total), po, lo, C.c_int64(slot0),
                                     C.c_int64(n), C.c_uint64(seed), C.c_uint64(gen),
                                     action.ctypes.data_as(C.POINTER(C.c_uint8)),
                                     loss.ctypes.data_as(C.POINTER(C.c_double)), C.byref(natt)))
    return action, loss, natt.value


def ww22_self_apply_fast(w, n_steps, n=None):
    n = w.shape[0] if n is None else n
    pw, ld = _f32(w)
    _chk(lib().srnn_oracle_ww22_self_apply_fast(pw, ld, C.c_int64(n), C.c_int32(n_steps)))


def num_threads():
    return int(lib().srnn_oracle_num_threads())


def set_threads(n):
    """OpenMP team size of the oracle (torchrun exports OMP_NUM_THREADS=1; the CPU baseline wants every core)."""
    lib().srnn_oracle_set_threads(C.c_int(int(n)))
