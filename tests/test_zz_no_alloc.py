"""GPU test of srnn_assign_uids_static -- the one entry point that takes no scratch from the caller.

It is meant for hosts that cannot allocate device memory themselves: bindings that only forward device pointers
owned by another runtime, and soup loops whose memory was handed out once at start-up.  The library keeps the
block-count scratch in a fixed array of its own (one per GPU), so the test checks two things a user relies on:
the result equals srnn_assign_uids' (exactly -- these are integers), and a second call right after the first is
not disturbed by what the first one left in that shared array.  Collected last on purpose (file name), and
self-contained so it also passes alone in a fresh process.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _both(eng, torch, dead_np, base):
    lib = eng.lib()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    n = dead_np.size
    dead = torch.from_numpy(dead_np).cuda()
    start = torch.arange(n, dtype=torch.int64, device="cuda")
    # existing entry point: caller-provided scratch
    uid_a = start.clone()
    tmp = torch.zeros((n + 1023) // 1024 + 1, dtype=torch.int64, device="cuda")
    nd_a = torch.zeros(1, dtype=torch.int64, device="cuda")
    rc = lib.srnn_assign_uids(C.c_void_p(dead.data_ptr()), C.c_int64(n), C.c_int64(base), C.c_void_p(uid_a.data_ptr()),
                              C.c_void_p(tmp.data_ptr()), C.c_void_p(nd_a.data_ptr()), stream)
    assert rc == 0, lib.srnn_last_error_string()
    # new entry point: no scratch argument at all
    uid_b = start.clone()
    nd_b = torch.zeros(1, dtype=torch.int64, device="cuda")
    rc = lib.srnn_assign_uids_static(C.c_void_p(dead.data_ptr()), C.c_int64(n), C.c_int64(base),
                                     C.c_void_p(uid_b.data_ptr()), C.c_void_p(nd_b.data_ptr()), stream)
    assert rc == 0, lib.srnn_last_error_string()
    torch.cuda.synchronize()
    return uid_a.cpu().numpy(), int(nd_a.item()), uid_b.cpu().numpy(), int(nd_b.item())


def test_static_scratch_matches_caller_scratch_twice_in_a_row():
    import torch
    import srnn_engine as eng
    torch.cuda.set_device(0)
    eng.lib().srnn_set_device(0)
    rng = np.random.default_rng(0)
    # first call: many blocks, dense deaths; second call: fewer blocks, sparse deaths -- stale counters from the
    # first call sit in the shared array while the second runs
    for n, rate, base in ((300_007, 0.6, 10 ** 9), (5_000, 0.01, 42), (1, 1.0, 7)):
        dead = (rng.random(n) < rate).astype(np.uint8)
        ua, na, ub, nb = _both(eng, torch, dead, base)
        assert na == nb == int(dead.sum())
        assert np.array_equal(ua, ub)
        want = np.arange(n)
        want[dead.astype(bool)] = base + np.arange(int(dead.sum()))
        assert np.array_equal(ub, want)


def test_static_scratch_refuses_oversize():
    import torch
    import srnn_engine as eng
    lib = eng.lib()
    dead = torch.zeros(8, dtype=torch.uint8, device="cuda")
    uid = torch.zeros(8, dtype=torch.int64, device="cuda")
    rc = lib.srnn_assign_uids_static(C.c_void_p(dead.data_ptr()), C.c_int64(eng.UIDS_STATIC_MAX + 1), C.c_int64(0),
                                     C.c_void_p(uid.data_ptr()), None, None)
    assert rc == 4                                           # SRNN_ERR_LIMIT, nothing launched
    assert b"SRNN_UIDS_STATIC_MAX" in lib.srnn_last_error_string()
    torch.cuda.synchronize()
    assert int(uid.abs().sum().item()) == 0
