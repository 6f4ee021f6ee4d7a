"""CPU tests of the drop-in boundary: the shared library loads, exports every symbol the header
declares, and every compute entry point refuses to run without a GPU (no silent fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT
from cslam_b200 import _native

HAS_GPU = torch.cuda.is_available()


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "cslam_nns.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cslam_nns_[a-z_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = _native.load()
    declared = _declared_symbols()
    assert len(declared) >= 18
    for name in declared:
        assert hasattr(lib, name), name
    assert set(declared) == set(_native.SIGNATURES) | set(_native.XCHG_SIGNATURES), \
        "ctypes table out of sync with the header"
    assert lib.cslam_nns_abi_version() == 1


def test_library_has_no_libcuda_link_dependency():
    # the driver API is reached through cudaGetDriverEntryPoint so the .so loads without a driver
    out = os.popen("ldd %s" % _native.LIB_PATH).read()
    assert "libcuda.so" not in out and "libtorch" not in out


@pytest.mark.skipif(HAS_GPU, reason="checks the no-GPU failure mode")
def test_exchange_create_fails_loudly_without_gpu():
    lib = _native.load()
    x = ctypes.c_void_p()
    assert lib.cslam_nns_xchg_create(0, 2, 0, ctypes.byref(x)) == _native.ENODEV
    assert not x.value
    assert lib.cslam_nns_xchg_create(0, 9, 0, ctypes.byref(x)) == _native.EINVAL   # world <= 8
    assert b"cslam_nns_xchg_create" in lib.cslam_nns_xchg_last_error(None)


@pytest.mark.skipif(HAS_GPU, reason="checks the no-GPU failure mode")
def test_create_fails_loudly_without_gpu():
    lib = _native.load()
    h = ctypes.c_void_p()
    rc = lib.cslam_nns_create(0, 16, 0, ctypes.byref(h))
    assert rc == _native.ENODEV
    assert not h.value
    assert b"no CPU fallback" in lib.cslam_nns_last_error(None)


@pytest.mark.skipif(HAS_GPU, reason="checks the no-GPU failure mode")
def test_python_class_fails_loudly_without_gpu():
    from cslam_b200 import NearestNeighborsMatching
    m = NearestNeighborsMatching()          # lazy: nothing allocated yet, like the reference
    assert m.search(np.ones(4), 1) == ([], [])
    assert m.search_best(np.ones(4)) == (None, None)
    with pytest.raises(Exception):
        m.add_item(np.ones(4), 0)           # first allocation needs the device
    with pytest.raises(Exception):
        NearestNeighborsMatching(dim=8)


def test_reference_api_surface():
    import inspect
    from cslam_b200 import NearestNeighborsMatching as M
    # signatures of cslam/nns_matching.py:9,23,44,65
    assert list(inspect.signature(M.__init__).parameters) == ["self", "dim"]
    assert inspect.signature(M.__init__).parameters["dim"].default is None
    assert list(inspect.signature(M.add_item).parameters) == ["self", "vector", "item"]
    assert list(inspect.signature(M.search).parameters) == ["self", "query", "k"]
    assert list(inspect.signature(M.search_best).parameters) == ["self", "query"]
    m = M()
    assert (m.n, m.dim, m.items, m.data) == (0, None, {}, [])
    assert m.device == torch.device("cuda")


def test_product_never_imports_oracle():
    for dirpath, _d, files in os.walk(os.path.join(ROOT, "cslam_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.lower().replace("test infrastructure", ""), f


def test_host_query_fast_path_rules():
    """Which queries may take the no-torch fast path of search / search_best_multi (host fp32, contiguous,
    exactly dim values) and that a wrong length raises RuntimeError like the reference's torch reshape
    (cslam/nns_matching.py:60)."""
    from cslam_b200 import NearestNeighborsMatching as M
    m = M()
    m.dim = 8
    q = np.arange(8, dtype=np.float32)
    keep, ptr = m._host_query(q)
    assert keep is q and ptr == q.ctypes.data
    strided = np.arange(16, dtype=np.float32)[::2]
    keep, ptr = m._host_query(strided)
    assert keep.flags.c_contiguous and np.array_equal(keep, strided) and ptr == keep.ctypes.data
    with pytest.raises(RuntimeError):
        m._host_query(np.zeros(7, np.float32))
    t = torch.arange(8, dtype=torch.float32)
    assert m._host_query(t) == (t, t.data_ptr()) or m._host_query(t)[1] == t.data_ptr()
    assert m._host_query(t.reshape(1, 8))[1] == t.data_ptr()
    assert m._host_query(t.double()) == (None, 0)              # float64 torch query: promotion path
    assert m._host_query(torch.arange(16, dtype=torch.float32)[::2]) == (None, 0)   # strided
    assert m._host_query(torch.zeros(7)) == (None, 0)          # wrong length: the general path raises


def test_integer_item_tracking_for_the_device_filter():
    """ids reach the device id table only while EVERY item is an int (keyframe ids); bools, strings and
    out-of-range ints switch the table off for good (search_candidates then refuses an id filter)."""
    from cslam_b200 import NearestNeighborsMatching as M
    m = M()
    m._note_ids([1, np.int64(2), np.int32(-3)])
    assert m._ids_ok and m._ids_pending == [1, 2, -3]
    m._note_ids([2 ** 63])
    assert not m._ids_ok and m._ids_pending == []
    for bad in (True, "kf", 1.5, None, (1, 2)):
        m2 = M()
        m2._note_ids([7, bad])
        assert not m2._ids_ok and m2._ids_pending == []
    assert not M().ids_on_device        # nothing allocated yet


def test_items_lookup_helpers_need_no_gpu():
    """Row -> item mapping of the batched calls (one C-level itemgetter call per block; identity for bulk shards)."""
    import numpy as np
    from cslam_b200 import NearestNeighborsMatching
    from cslam_b200.sharded import RowIdentity, ShardedNearestNeighborsMatching
    m = object.__new__(NearestNeighborsMatching)   # the constructor needs a CUDA device (it fails loudly without)
    m.items = {0: "a", 1: "b", 2: ("c", 3), 3: None}
    assert m._items_of(np.array([[0, 1], [3, 2]])) == [["a", "b"], [None, ("c", 3)]]
    assert m._items_of(np.array([[2]])) == [[("c", 3)]]
    assert m._items_of(np.empty((3, 0), np.int64)) == [[], [], []]
    s = object.__new__(ShardedNearestNeighborsMatching)
    s.items = RowIdentity()
    assert s._items_of(np.array([[5, 7], [9, 11]])) == [[5, 7], [9, 11]]
    s.items = {5: "x", 7: "y"}
    assert s._items_of(np.array([[7, 5]])) == [["y", "x"]]
