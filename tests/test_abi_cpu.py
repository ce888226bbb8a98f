"""CPU tests of the drop-in boundary: the C-ABI library loads without a GPU and exports every symbol that
include/srnn_b200.h declares; host-only entry points (descriptor maths) agree with the oracle."""
import ctypes as C
import os
import re

import pytest

import oracle
import srnn_engine as eng

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "srnn_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(srnn_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    lib = eng.lib()
    names = _declared()
    assert len(names) >= 15
    for name in names:
        assert hasattr(lib, name), "libsrnn_b200.so does not export %s" % name
    assert sorted(eng.EXPORTS) == names


def test_version_and_error_string():
    lib = eng.lib()
    assert lib.srnn_version() >= 100
    assert isinstance(lib.srnn_last_error_string(), bytes)


@pytest.mark.parametrize("kind,width,depth,aggs,nw", [(0, 2, 2, 0, 14), (1, 2, 2, 4, 20), (2, 2, 2, 0, 17),
                                                       (0, 3, 2, 0, 24), (1, 2, 2, 3, 16), (2, 2, 1, 0, 9)])
def test_desc_init_matches_oracle(kind, width, depth, aggs, nw):
    d = eng.make_desc(kind, width, depth, aggs)
    assert d.n_weights == nw == oracle.desc(kind, width, depth, aggs).n_weights
    assert sum(a * b for a, b in eng.layer_shapes(d)) == nw
    assert eng.shape_supported(d)


def test_bad_shapes_are_errors_not_crashes():
    with pytest.raises(eng.SrnnError):
        eng.make_desc(7, 2, 2)
    with pytest.raises(eng.SrnnError):
        eng.make_desc(0, 99, 2)
    with pytest.raises(eng.SrnnError):
        eng.make_desc(1, 1, 3, 4)          # 10 weights, 4 aggregates: leftovers (2) fill another collection
    assert eng.shape_supported(eng.make_desc(0, 4, 3)) == 2     # valid, served by the generic run-time-shape path
    assert eng.shape_supported(eng.make_desc(0, 2, 2)) == 1


def test_comm_family_argument_checks():
    """srnn_comm_*: argument errors are status codes (checked before NCCL is even looked up); destroying NULL is a no-op."""
    lib = eng.lib()
    assert lib.srnn_comm_unique_id(None) == 1 and b"NULL" in lib.srnn_last_error_string()
    comm = C.c_void_p()
    assert lib.srnn_comm_init(None, C.c_int32(0), C.c_int32(1), C.byref(comm)) == 1
    uid = (C.c_uint8 * 128)()
    assert lib.srnn_comm_init(uid, C.c_int32(2), C.c_int32(2), C.byref(comm)) == 1      # rank out of range
    assert b"rank" in lib.srnn_last_error_string()
    assert lib.srnn_comm_allgather(None, None, None, C.c_int64(4), None) == 1
    assert lib.srnn_comm_destroy(None) == 0


def test_no_cpu_fallback():
    """Without a CUDA device the product path must refuse, not fall back."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(eng.SrnnError):
        eng.Population(eng.make_desc(0, 2, 2), 16)
