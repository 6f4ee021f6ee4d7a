"""ctypes binding of ``libcslam_nns.so`` (the C-ABI declared in ``include/cslam_nns.h``).

The library is built in-tree by ``make -C cslam_b200/csrc`` (``__graft_entry__.build()`` does it)
and lives at ``cslam_b200/_lib/libcslam_nns.so``.  There is deliberately no fallback: if the
library is missing, or no sm_100 device is usable, the error is raised to the caller.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_lib", "libcslam_nns.so")

OK = 0
EINVAL, ENODEV, ENOMEM, ECUDA, ESTATE = -1, -2, -3, -4, -5
FILTER_BEST, FILTER_INTRA = 0, 1  # CSLAM_NNS_FILTER_*

_c = ctypes
_handle = _c.c_void_p
_i64 = _c.c_int64
_f32p = _c.c_void_p  # raw addresses (host numpy / device pointers) are passed as integers
_i64p = _c.c_void_p

# name -> (restype, argtypes); mirrors include/cslam_nns.h one to one
SIGNATURES = {
    "cslam_nns_abi_version": (_c.c_int, []),
    "cslam_nns_launch_count": (_i64, []),
    "cslam_nns_last_error": (_c.c_char_p, [_handle]),
    "cslam_nns_create": (_c.c_int, [_c.c_int, _i64, _i64, _c.POINTER(_handle)]),
    "cslam_nns_destroy": (_c.c_int, [_handle]),
    "cslam_nns_set_dim": (_c.c_int, [_handle, _i64]),
    "cslam_nns_reserve": (_c.c_int, [_handle, _i64]),
    "cslam_nns_size": (_i64, [_handle]),
    "cslam_nns_capacity": (_i64, [_handle]),
    "cslam_nns_dim": (_i64, [_handle]),
    "cslam_nns_ld": (_i64, [_handle]),
    "cslam_nns_data_ptr": (_c.c_void_p, [_handle]),
    "cslam_nns_device": (_c.c_int, [_handle]),
    "cslam_nns_add": (_c.c_int, [_handle, _f32p, _i64, _i64, _c.c_int, _c.c_void_p,
                                 _c.POINTER(_i64)]),
    "cslam_nns_search": (_c.c_int, [_handle, _f32p, _i64, _i64, _c.c_int, _c.c_int, _i64p, _f32p,
                                    _c.c_int, _c.c_void_p, _c.POINTER(_i64)]),
    "cslam_nns_search_multi": (_c.c_int, [_c.POINTER(_handle), _c.c_int, _f32p, _i64, _c.c_int, _c.c_int,
                                          _i64p, _f32p, _i64p, _c.c_void_p]),
    "cslam_nns_set_row_ids": (_c.c_int, [_handle, _i64, _i64, _i64p]),
    "cslam_nns_search_filtered": (_c.c_int, [_handle, _f32p, _i64, _i64, _c.c_int, _c.c_int, _c.c_int,
                                             _i64p, _i64, _c.c_float, _i64p, _f32p, _c.c_void_p,
                                             _c.c_void_p, _c.POINTER(_i64)]),
    "cslam_nns_search_keys": (_c.c_int, [_handle, _f32p, _i64, _i64, _c.c_int, _c.c_int, _i64, _i64,
                                         _c.c_void_p, _c.c_void_p]),
    "cslam_nns_merge_keys": (_c.c_int, [_c.c_int, _c.c_void_p, _c.c_int, _i64, _c.c_int, _i64p,
                                        _f32p, _c.c_int, _c.c_void_p, _c.POINTER(_i64)]),
    "cslam_nns_profile": (_c.c_int, [_handle, _c.c_int]),
    "cslam_nns_profile_read": (_c.c_int, [_handle, _c.POINTER(_c.c_float), _c.POINTER(_i64)]),
    "cslam_nns_set_variant": (_c.c_int, [_handle, _c.c_int]),
    "cslam_nns_last_variant": (_c.c_int, [_handle]),
    "cslam_nns_set_screen": (_c.c_int, [_handle, _c.c_int, _i64]),
    "cslam_nns_screen_stats": (_c.c_int, [_handle, _c.POINTER(_i64), _c.POINTER(_i64)]),
    "cslam_nns_debug_copy": (_c.c_int, [_handle, _c.c_int, _c.c_void_p, _i64]),
    "cslam_nns_search_static": (_c.c_int, [_c.c_int, _f32p, _i64, _i64, _i64, _f32p, _c.c_int, _i64p, _f32p,
                                            _c.c_void_p, _c.c_char_p, _i64]),
    "cslam_nns_time_search": (_c.c_int, [_handle, _f32p, _i64, _i64, _c.c_int, _c.c_int,
                                         _c.c_void_p, _c.POINTER(_c.c_float)]),
}

# include/cslam_scancontext.h (ScanContext matcher)
_f64p = _c.c_void_p
XCHG_SIGNATURES = {
    "cslam_nns_xchg_last_error": (_c.c_char_p, [_handle]),
    "cslam_nns_xchg_create": (_c.c_int, [_c.c_int, _c.c_int, _c.c_int, _c.POINTER(_handle)]),
    "cslam_nns_xchg_destroy": (_c.c_int, [_handle]),
    "cslam_nns_xchg_export": (_c.c_int, [_handle, _c.c_void_p]),
    "cslam_nns_xchg_connect": (_c.c_int, [_handle, _c.c_void_p]),
    "cslam_nns_xchg_connect_local": (_c.c_int, [_handle, _c.POINTER(_handle)]),
    "cslam_nns_xchg_merge": (_c.c_int, [_handle, _c.c_void_p, _i64, _c.c_int, _c.c_int, _c.c_void_p,
                                        _c.c_void_p, _c.c_void_p]),
    "cslam_nns_xchg_status": (_c.c_int, [_handle]),
    "cslam_nns_search_sharded": (_c.c_int, [_handle, _handle, _f32p, _i64, _i64, _c.c_int, _c.c_int, _c.c_int,
                                            _c.c_int, _c.c_int, _c.c_void_p, _c.c_void_p, _c.c_void_p]),
}

SC_SIGNATURES = {
    "cslam_sc_last_error": (_c.c_char_p, [_handle]),
    "cslam_sc_create": (_c.c_int, [_c.c_int, _c.c_int, _c.c_int, _c.c_int, _i64, _c.POINTER(_handle)]),
    "cslam_sc_destroy": (_c.c_int, [_handle]),
    "cslam_sc_size": (_i64, [_handle]),
    "cslam_sc_add": (_c.c_int, [_handle, _f64p, _f64p, _c.c_void_p, _c.POINTER(_i64)]),
    "cslam_sc_search": (_c.c_int, [_handle, _f64p, _f64p, _i64p, _f64p, _c.c_void_p, _c.c_void_p]),
    "cslam_sc_descriptors_ptr": (_c.c_void_p, [_handle]),
    "cslam_sc_ringkeys_ptr": (_c.c_void_p, [_handle]),
}

_lib = None


class NativeError(RuntimeError):
    """A C-ABI call failed; ``code`` is the CSLAM_NNS_E* value."""

    def __init__(self, code, message):
        super().__init__(f"cslam_nns error {code}: {message}")
        self.code = code


def load():
    """Load the shared library (once).  Raises ``ImportError`` if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `make -C cslam_b200/csrc` "
            "(or `python -c 'import __graft_entry__ as g; g.build()'`). "
            "cslam_b200 has no CPU or PyTorch fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in list(SIGNATURES.items()) + list(SC_SIGNATURES.items()) + list(XCHG_SIGNATURES.items()):
        fn = getattr(lib, name)  # AttributeError here == header/library mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, handle=None):
    if rc == OK:
        return
    msg = load().cslam_nns_last_error(handle)
    raise NativeError(rc, msg.decode("utf-8", "replace") if msg else "")
