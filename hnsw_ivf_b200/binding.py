"""ctypes binding of include/b200ivf_c.h (the drop-in C-ABI).

Buffers may be numpy arrays (host) or torch tensors (host or CUDA); CUDA tensors are passed as raw
device pointers, so a search on resident inputs performs no host<->device copy.
"""
import ctypes as C
import os
import subprocess

import numpy as np

METRIC_INNER_PRODUCT = 0  # faiss/MetricType.h:23
METRIC_L2 = 1             # faiss/MetricType.h:24

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class B200Error(RuntimeError):
    """Non-zero return code of the C-ABI (c_api/error_c.h:19-30 codes)."""

    def __init__(self, code, msg):
        super().__init__("b200ivf error %d: %s" % (code, msg))
        self.code = code


class IndexIVFStats(C.Structure):  # FaissIndexIVFStats, c_api/IndexIVF_c.h:162-169
    _fields_ = [("nq", C.c_size_t), ("nlist", C.c_size_t), ("ndis", C.c_size_t),
                ("nheap_updates", C.c_size_t), ("quantization_time", C.c_double),
                ("search_time", C.c_double)]


class QueryLatencyStats(C.Structure):  # the fork's faiss/IndexIVF.h:28-32, microseconds
    _fields_ = [("total_us", C.c_double), ("quantization_us", C.c_double),
                ("list_scan_us", C.c_double)]


def library_path():
    return os.path.join(_HERE, "lib", "libb200ivf.so")


def build_library(verbose=False):
    """Compile csrc/*.cu for sm_100a into lib/libb200ivf.so (nvcc cross-compiles without a GPU)."""
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-j8", "-C", os.path.join(_HERE, "csrc")], stdout=out)
    return library_path()


# every symbol include/b200ivf_c.h declares: (name, restype, argtypes)
_vp, _f32p, _i64p = C.c_void_p, C.c_void_p, C.c_void_p
SYMBOLS = [
    ("b200_get_last_error", C.c_char_p, []),
    ("b200_IndexIVFFlat_new_with_metric", C.c_int,
     [C.POINTER(_vp), C.c_size_t, C.c_size_t, C.c_int, C.c_int]),
    ("b200_IndexIVFFlat_free", None, [_vp]),
    ("b200_IndexIVFFlat_nlist", C.c_size_t, [_vp]),
    ("b200_IndexIVFFlat_nprobe", C.c_size_t, [_vp]),
    ("b200_IndexIVFFlat_set_nprobe", None, [_vp, C.c_size_t]),
    ("b200_IndexIVFFlat_d", C.c_int, [_vp]),
    ("b200_IndexIVFFlat_ntotal", C.c_int64, [_vp]),
    ("b200_IndexIVFFlat_is_trained", C.c_int, [_vp]),
    ("b200_IndexIVFFlat_metric_type", C.c_int, [_vp]),
    ("b200_IndexIVFFlat_set_centroids", C.c_int, [_vp, _f32p]),
    ("b200_IndexIVFFlat_get_centroids", C.c_int, [_vp, _f32p]),
    ("b200_IndexIVFFlat_set_list_range", C.c_int, [_vp, C.c_size_t, C.c_size_t]),
    ("b200_IndexIVFFlat_add_with_ids", C.c_int, [_vp, C.c_int64, _f32p, _i64p]),
    ("b200_IndexIVFFlat_add_core", C.c_int, [_vp, C.c_int64, _f32p, _i64p, _i64p]),
    ("b200_IndexIVFFlat_reset", C.c_int, [_vp]),
    ("b200_IndexIVFFlat_get_list_size", C.c_size_t, [_vp, C.c_size_t]),
    ("b200_IndexIVFFlat_invlists_get_ids", C.c_int, [_vp, C.c_size_t, _i64p]),
    ("b200_IndexIVFFlat_invlists_get_codes", C.c_int, [_vp, C.c_size_t, _f32p]),
    ("b200_IndexIVFFlat_search", C.c_int, [_vp, C.c_int64, _f32p, C.c_int64, _f32p, _i64p]),
    ("b200_IndexIVFFlat_search_preassigned", C.c_int,
     [_vp, C.c_int64, _f32p, C.c_int64, _i64p, _f32p, _f32p, _i64p, C.c_int]),
    ("b200_IndexIVFFlat_quantize", C.c_int, [_vp, C.c_int64, _f32p, C.c_int64, _f32p, _i64p]),
    ("b200_IndexIVFFlat_quantize_noalloc", C.c_int,
     [_vp, C.c_int64, _f32p, C.c_int64, _f32p, _i64p]),
    ("b200_IndexIVFFlat_get_stats", None, [_vp, C.POINTER(IndexIVFStats)]),
    ("b200_IndexIVFFlat_reset_stats", None, [_vp]),
    ("b200_IndexIVFFlat_last_timings", None, [_vp, C.POINTER(C.c_double)]),
    ("b200_IndexIVFFlat_set_stream", C.c_int, [_vp, _vp]),
    ("b200_debug_trap_info", C.c_int, [C.POINTER(C.c_ulonglong)]),
    ("b200_IndexIVFFlat_set_option", C.c_int, [_vp, C.c_char_p, C.c_double]),
    ("b200_kernel_launches", C.c_ulonglong, []),
    ("b200_tensor_core_launches", C.c_ulonglong, []),
    ("b200_merge_knn_results", C.c_int,
     [C.c_int, C.c_int, C.c_int64, C.c_int64, _f32p, _i64p, _f32p, _i64p, C.c_int, _vp]),
    ("b200_IndexIVFFlat_train", C.c_int, [_vp, C.c_int64, _f32p]),
    ("b200_IDSelectorRange_new", C.c_int, [C.POINTER(_vp), C.c_int64, C.c_int64]),
    ("b200_IDSelectorBatch_new", C.c_int, [C.POINTER(_vp), C.c_size_t, _i64p, C.c_int]),
    ("b200_IDSelectorBitmap_new", C.c_int, [C.POINTER(_vp), C.c_size_t, _vp, C.c_int]),
    ("b200_IDSelectorNot_new", C.c_int, [C.POINTER(_vp), _vp]),
    ("b200_IDSelectorAnd_new", C.c_int, [C.POINTER(_vp), _vp, _vp]),
    ("b200_IDSelectorOr_new", C.c_int, [C.POINTER(_vp), _vp, _vp]),
    ("b200_IDSelectorXOr_new", C.c_int, [C.POINTER(_vp), _vp, _vp]),
    ("b200_IDSelector_is_member", C.c_int, [_vp, C.c_int64]),
    ("b200_IDSelector_free", None, [_vp]),
    ("b200_SearchParametersIVF_new", C.c_int, [C.POINTER(_vp)]),
    ("b200_SearchParametersIVF_new_with", C.c_int, [C.POINTER(_vp), _vp, C.c_size_t, C.c_size_t]),
    ("b200_SearchParametersIVF_free", None, [_vp]),
    ("b200_SearchParametersIVF_sel", _vp, [_vp]),
    ("b200_SearchParametersIVF_nprobe", C.c_size_t, [_vp]),
    ("b200_SearchParametersIVF_set_nprobe", None, [_vp, C.c_size_t]),
    ("b200_SearchParametersIVF_max_codes", C.c_size_t, [_vp]),
    ("b200_SearchParametersIVF_set_max_codes", None, [_vp, C.c_size_t]),
    ("b200_IndexIVFFlat_max_codes", C.c_size_t, [_vp]),
    ("b200_IndexIVFFlat_set_max_codes", None, [_vp, C.c_size_t]),
    ("b200_IndexIVFFlat_search_with_params", C.c_int,
     [_vp, C.c_int64, _f32p, C.c_int64, _vp, _f32p, _i64p]),
    ("b200_IndexIVFFlat_search_preassigned_with_params", C.c_int,
     [_vp, C.c_int64, _f32p, C.c_int64, _i64p, _f32p, _f32p, _i64p, C.c_int, _vp]),
    ("b200_IndexIVFFlat_search_stats", C.c_int,
     [_vp, C.c_int64, _f32p, C.c_int64, _f32p, _i64p, _vp, C.POINTER(QueryLatencyStats)]),
    ("b200_RangeSearchResult_new", C.c_int, [C.POINTER(_vp), C.c_int64]),
    ("b200_RangeSearchResult_free", None, [_vp]),
    ("b200_RangeSearchResult_nq", C.c_size_t, [_vp]),
    ("b200_RangeSearchResult_lims", None, [_vp, C.POINTER(C.POINTER(C.c_size_t))]),
    ("b200_RangeSearchResult_labels", None,
     [_vp, C.POINTER(C.POINTER(C.c_int64)), C.POINTER(C.POINTER(C.c_float))]),
    ("b200_IndexIVFFlat_range_search", C.c_int, [_vp, C.c_int64, _f32p, C.c_float, _vp, _vp]),
    ("b200_IndexIVFFlat_range_search_preassigned", C.c_int,
     [_vp, C.c_int64, _f32p, C.c_float, _i64p, _f32p, _vp, C.c_int, _vp]),
    ("b200_IndexIVFFlat_reserve_lists", C.c_int, [_vp, _i64p]),
    ("b200_IndexIVFFlat_get_list_sizes", C.c_int, [_vp, _i64p]),
    ("b200_IndexIVFFlat_stage_time_sums", C.c_ulonglong, [_vp, C.POINTER(C.c_double)]),
    ("b200_IndexIVFFlat_exact_redo_count", C.c_ulonglong, [_vp]),
    ("b200_IndexIVFFlat_synchronize", C.c_int, [_vp]),
    ("b200_nccl_get_unique_id", C.c_int, [_vp]),
    ("b200_IndexShardsIVFFlat_new", C.c_int,
     [C.POINTER(_vp), C.c_size_t, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int, _vp]),
    ("b200_IndexShardsIVFFlat_free", None, [_vp]),
    ("b200_IndexShardsIVFFlat_local", _vp, [_vp]),
    ("b200_IndexShardsIVFFlat_rank", C.c_int, [_vp]),
    ("b200_IndexShardsIVFFlat_world", C.c_int, [_vp]),
    ("b200_IndexShardsIVFFlat_ntotal", C.c_int64, [_vp]),
    ("b200_IndexShardsIVFFlat_set_centroids", C.c_int, [_vp, _f32p, C.c_int]),
    ("b200_IndexShardsIVFFlat_train", C.c_int, [_vp, C.c_int64, _f32p, C.c_int]),
    ("b200_IndexShardsIVFFlat_add_with_ids", C.c_int, [_vp, C.c_int64, _f32p, _i64p]),
    ("b200_IndexShardsIVFFlat_add_core", C.c_int, [_vp, C.c_int64, _f32p, _i64p, _i64p]),
    ("b200_IndexShardsIVFFlat_search", C.c_int,
     [_vp, C.c_int64, _f32p, C.c_int64, _f32p, _i64p, C.c_int]),
    ("b200_IndexShardsIVFFlat_phase_time_sums", C.c_ulonglong, [_vp, C.POINTER(C.c_double)]),
    ("b200_IndexShardsIVFFlat_reset_phase_times", None, [_vp]),
    ("b200_IndexShardsIVFFlat_synchronize", C.c_int, [_vp]),
    ("b200_write_index_fname", C.c_int, [_vp, C.c_char_p]),
    ("b200_read_index_fname", C.c_int, [C.c_char_p, C.c_int, C.c_int, C.POINTER(_vp)]),
]


def load_library():
    """dlopen the C-ABI library and bind every declared symbol. Raises if it is not built."""
    global _LIB
    if _LIB is None:
        path = library_path()
        if not os.path.exists(path):
            raise B200Error(-2, "%s is not built (run __graft_entry__.build()); "
                            "there is no CPU fallback" % path)
        lib = C.CDLL(path)
        for name, res, args in SYMBOLS:
            fn = getattr(lib, name)  # AttributeError if the .so lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _LIB = lib
    return _LIB


# ------------------------------------------------------------------------------------------------
def _is_torch(a):
    return type(a).__module__.startswith("torch")


def _ptr(a, dtype, shape=None, name="buffer"):
    """(keepalive, raw pointer) of a numpy array or torch tensor, checked for dtype/contiguity."""
    if a is None:
        return None, None
    if _is_torch(a):
        import torch
        want = {np.float32: torch.float32, np.int64: torch.int64}[dtype]
        if a.dtype != want:
            raise TypeError("%s must be %s" % (name, want))
        if not a.is_contiguous():
            raise ValueError("%s must be contiguous" % name)
        if shape is not None and tuple(a.shape) != tuple(shape):
            raise ValueError("%s has shape %s, expected %s" % (name, tuple(a.shape), shape))
        return a, a.data_ptr()
    arr = np.ascontiguousarray(a, dtype=dtype)
    if shape is not None and tuple(arr.shape) != tuple(shape):
        raise ValueError("%s has shape %s, expected %s" % (name, arr.shape, shape))
    return arr, arr.ctypes.data


def _empty_like_input(x, shape, dtype):
    """Result buffer living where the queries live (CUDA tensor -> CUDA tensor, else numpy)."""
    if _is_torch(x):
        import torch
        td = {np.float32: torch.float32, np.int64: torch.int64}[dtype]
        return torch.empty(shape, dtype=td, device=x.device)
    return np.empty(shape, dtype=dtype)


def _rc(lib, rc):
    if rc != 0:
        raise B200Error(rc, lib.b200_get_last_error().decode())


class IDSelector:
    """Base of the selector family (faiss/impl/IDSelector.h:19-170): owns a C-ABI handle and keeps
    its operands alive (the C objects borrow them, as the reference's do)."""

    def __init__(self, handle, keep=()):
        self._lib = load_library()
        self._h = handle
        self._keep = keep

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.b200_IDSelector_free(self._h)
            self._h = None

    def is_member(self, i):
        return bool(self._lib.b200_IDSelector_is_member(self._h, int(i)))


def _new_sel(fn_name, *args, keep=()):
    lib = load_library()
    h = _vp()
    _rc(lib, getattr(lib, fn_name)(C.byref(h), *args))
    return IDSelector(h, keep)


def IDSelectorRange(imin, imax):
    """ids in [imin, imax) (faiss/impl/IDSelector.cpp:20-22)"""
    return _new_sel("b200_IDSelectorRange_new", int(imin), int(imax))


def IDSelectorBatch(indices, device=0):
    """ids in a set (faiss/impl/IDSelector.cpp:84-108); IDSelectorArray is the same predicate"""
    a = np.ascontiguousarray(indices, dtype=np.int64)
    return _new_sel("b200_IDSelectorBatch_new", a.size, a.ctypes.data, device)


IDSelectorArray = IDSelectorBatch


def IDSelectorBitmap(bitmap, device=0):
    """bit i of a uint8 bitmap, ids beyond it excluded (faiss/impl/IDSelector.cpp:117-123)"""
    a = np.ascontiguousarray(bitmap, dtype=np.uint8)
    return _new_sel("b200_IDSelectorBitmap_new", a.size, a.ctypes.data, device)


def IDSelectorNot(sel):
    return _new_sel("b200_IDSelectorNot_new", sel._h, keep=(sel,))


def IDSelectorAnd(lhs, rhs):
    return _new_sel("b200_IDSelectorAnd_new", lhs._h, rhs._h, keep=(lhs, rhs))


def IDSelectorOr(lhs, rhs):
    return _new_sel("b200_IDSelectorOr_new", lhs._h, rhs._h, keep=(lhs, rhs))


def IDSelectorXOr(lhs, rhs):
    return _new_sel("b200_IDSelectorXOr_new", lhs._h, rhs._h, keep=(lhs, rhs))


class SearchParametersIVF:
    """faiss.SearchParametersIVF(sel=None, nprobe=1, max_codes=0) (faiss/IndexIVF.h:73-88)"""

    def __init__(self, sel=None, nprobe=1, max_codes=0):
        self._lib = load_library()
        self._sel = sel
        h = _vp()
        _rc(self._lib, self._lib.b200_SearchParametersIVF_new_with(
            C.byref(h), sel._h if sel is not None else None, int(nprobe), int(max_codes)))
        self._h = h

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.b200_SearchParametersIVF_free(self._h)
            self._h = None

    sel = property(lambda s: s._sel)

    @property
    def nprobe(self):
        return self._lib.b200_SearchParametersIVF_nprobe(self._h)

    @nprobe.setter
    def nprobe(self, v):
        self._lib.b200_SearchParametersIVF_set_nprobe(self._h, int(v))

    @property
    def max_codes(self):
        return self._lib.b200_SearchParametersIVF_max_codes(self._h)

    @max_codes.setter
    def max_codes(self, v):
        self._lib.b200_SearchParametersIVF_set_max_codes(self._h, int(v))


class IndexIVFFlat:
    """Device-resident IndexIVFFlat with its flat quantizer (faiss/IndexIVFFlat.h:26).

    Same field names and call shapes as the reference's Python surface: ``d``, ``nlist``,
    ``nprobe``, ``ntotal``, ``is_trained``, ``metric_type``; ``add``, ``add_with_ids``,
    ``max_codes``; ``train``, ``add``, ``add_with_ids``, ``search(x, k, params=None) -> (D, I)``,
    ``search_preassigned``, ``range_search(x, radius) -> (lims, D, I)``, and the fork's
    ``search_stats``. Centroids trained elsewhere can be installed with ``set_centroids``.
    """

    def __init__(self, d, nlist, metric=METRIC_L2, device=0, _handle=None, _borrowed=False):
        self._lib = load_library()
        h = _vp()
        self._h = None
        if _handle is not None:
            h = _handle
        else:
            self._check(self._lib.b200_IndexIVFFlat_new_with_metric(C.byref(h), d, nlist, metric,
                                                                    device))
        self._h = h
        self._borrowed = _borrowed  # a shard of an IndexShardsIVFFlat: freed with its owner
        self.device = device
        self._own_stream = True

    def __del__(self):
        if getattr(self, "_h", None) and not getattr(self, "_borrowed", False):
            self._lib.b200_IndexIVFFlat_free(self._h)
        self._h = None

    def _settle(self, *arrays):
        """All-device calls return once their work is enqueued on the handle's stream
        (b200ivf_c.h, set_stream). While that is the handle's own stream nothing orders torch's
        next use of the results with it, so wait here; after ``set_stream(torch's stream)`` the
        call stays asynchronous and stream-ordered."""
        if self._own_stream and any(a is not None and _is_torch(a) and a.is_cuda for a in arrays):
            self._check(self._lib.b200_IndexIVFFlat_synchronize(self._h))

    def synchronize(self):
        """wait for asynchronous calls in flight; raises their deferred failure, if any"""
        self._check(self._lib.b200_IndexIVFFlat_synchronize(self._h))

    def _order_with_torch(self, *arrays):
        """While the handle runs on its own (non-blocking) stream it is not ordered with torch's:
        CUDA tensors handed to it may still be being produced there, and result tensors fresh from
        torch's caching allocator may still be in use by queued torch work. So: wait for torch's
        current stream first. After ``set_stream(torch.cuda.current_stream().cuda_stream)`` the
        library is stream-ordered with torch and nothing is waited for."""
        if self._own_stream and any(a is not None and _is_torch(a) and a.is_cuda for a in arrays):
            import torch
            torch.cuda.current_stream(self.device).synchronize()

    def _check(self, rc):
        if rc != 0:
            raise B200Error(rc, self._lib.b200_get_last_error().decode())

    # ---- fields ------------------------------------------------------------------------------
    d = property(lambda s: s._lib.b200_IndexIVFFlat_d(s._h))
    nlist = property(lambda s: s._lib.b200_IndexIVFFlat_nlist(s._h))
    ntotal = property(lambda s: s._lib.b200_IndexIVFFlat_ntotal(s._h))
    is_trained = property(lambda s: bool(s._lib.b200_IndexIVFFlat_is_trained(s._h)))
    metric_type = property(lambda s: s._lib.b200_IndexIVFFlat_metric_type(s._h))

    @property
    def nprobe(self):
        return self._lib.b200_IndexIVFFlat_nprobe(self._h)

    @nprobe.setter
    def nprobe(self, v):
        self._lib.b200_IndexIVFFlat_set_nprobe(self._h, int(v))

    @property
    def max_codes(self):
        return self._lib.b200_IndexIVFFlat_max_codes(self._h)

    @max_codes.setter
    def max_codes(self, v):
        self._lib.b200_IndexIVFFlat_set_max_codes(self._h, int(v))

    # ---- build -------------------------------------------------------------------------------
    def train(self, x):
        """IndexIVF::train (faiss/IndexIVF.cpp:1404): k-means of the coarse quantizer on device"""
        kx, px = _ptr(x, np.float32, None, "x")
        if kx.ndim != 2 or kx.shape[1] != self.d:
            raise ValueError("x must be [n][%d]" % self.d)
        self._order_with_torch(kx)
        self._check(self._lib.b200_IndexIVFFlat_train(self._h, kx.shape[0], px))

    def set_centroids(self, c):
        keep, p = _ptr(c, np.float32, (self.nlist, self.d), "centroids")
        self._order_with_torch(keep)
        self._check(self._lib.b200_IndexIVFFlat_set_centroids(self._h, p))

    def get_centroids(self):
        out = np.empty((self.nlist, self.d), np.float32)
        self._check(self._lib.b200_IndexIVFFlat_get_centroids(self._h, out.ctypes.data))
        return out

    def set_list_range(self, lo, hi):
        self._check(self._lib.b200_IndexIVFFlat_set_list_range(self._h, lo, hi))

    def add(self, x):
        self.add_with_ids(x, None)

    def add_with_ids(self, x, ids):
        kx, px = _ptr(x, np.float32, None, "x")
        n = kx.shape[0]
        ki, pi = _ptr(ids, np.int64, (n,), "ids")
        self._order_with_torch(kx, ki)
        self._check(self._lib.b200_IndexIVFFlat_add_with_ids(self._h, n, px, pi))

    def add_core(self, x, ids, assign):
        kx, px = _ptr(x, np.float32, None, "x")
        n = kx.shape[0]
        ki, pi = _ptr(ids, np.int64, (n,), "ids")
        ka, pa = _ptr(assign, np.int64, (n,), "assign")
        self._order_with_torch(kx, ki, ka)
        self._check(self._lib.b200_IndexIVFFlat_add_core(self._h, n, px, pi, pa))

    def reset(self):
        self._check(self._lib.b200_IndexIVFFlat_reset(self._h))

    def reserve_lists(self, sizes):
        """room for sizes[l] rows in list l ([nlist] int64) so that later adds append in place"""
        ks, ps = _ptr(sizes, np.int64, (self.nlist,), "sizes")
        self._order_with_torch(ks)
        self._check(self._lib.b200_IndexIVFFlat_reserve_lists(self._h, ps))

    def list_sizes(self):
        out = np.empty((self.nlist,), np.int64)
        self._check(self._lib.b200_IndexIVFFlat_get_list_sizes(self._h, out.ctypes.data))
        return out

    def get_list_size(self, l):
        return self._lib.b200_IndexIVFFlat_get_list_size(self._h, l)

    def get_list(self, l):
        n = self.get_list_size(l)
        codes = np.empty((n, self.d), np.float32)
        ids = np.empty((n,), np.int64)
        if n:
            self._check(self._lib.b200_IndexIVFFlat_invlists_get_codes(self._h, l,
                                                                       codes.ctypes.data))
            self._check(self._lib.b200_IndexIVFFlat_invlists_get_ids(self._h, l, ids.ctypes.data))
        return codes, ids

    # ---- search ------------------------------------------------------------------------------
    def search(self, x, k, D=None, I=None, params=None):
        kx, px = _ptr(x, np.float32, None, "x")
        n = kx.shape[0]
        if kx.ndim != 2 or kx.shape[1] != self.d:
            raise ValueError("x must be [n][%d]" % self.d)
        if D is None:
            D = _empty_like_input(kx, (n, k), np.float32)
        if I is None:
            I = _empty_like_input(kx, (n, k), np.int64)
        kd, pd = _ptr(D, np.float32, (n, k), "D")
        ki, pi = _ptr(I, np.int64, (n, k), "I")
        self._order_with_torch(kx, kd, ki)
        if params is None:
            self._check(self._lib.b200_IndexIVFFlat_search(self._h, n, px, k, pd, pi))
        else:
            self._check(self._lib.b200_IndexIVFFlat_search_with_params(self._h, n, px, k,
                                                                       params._h, pd, pi))
        self._settle(kx, kd, ki)
        return kd, ki

    def search_preassigned(self, x, k, assign, centroid_dis=None, D=None, I=None,
                           store_pairs=False, params=None):
        kx, px = _ptr(x, np.float32, None, "x")
        n = kx.shape[0]
        nprobe = min(params.nprobe if params is not None else self.nprobe, self.nlist)
        ka, pa = _ptr(assign, np.int64, (n, nprobe), "assign")
        kc, pc = _ptr(centroid_dis, np.float32, None, "centroid_dis")
        if D is None:
            D = _empty_like_input(kx, (n, k), np.float32)
        if I is None:
            I = _empty_like_input(kx, (n, k), np.int64)
        kd, pd = _ptr(D, np.float32, (n, k), "D")
        ki, pi = _ptr(I, np.int64, (n, k), "I")
        self._order_with_torch(kx, ka, kd, ki)
        if params is None and not store_pairs:
            self._check(self._lib.b200_IndexIVFFlat_search_preassigned(self._h, n, px, k, pa, pc,
                                                                       pd, pi, 0))
        else:
            self._check(self._lib.b200_IndexIVFFlat_search_preassigned_with_params(
                self._h, n, px, k, pa, pc, pd, pi, int(bool(store_pairs)),
                params._h if params is not None else None))
        self._settle(kx, kd, ki)
        return kd, ki

    def search_stats(self, x, k, params=None):
        """The fork's IndexIVF::search_stats (faiss/IndexIVF.cpp:727): (D, I, stats) with stats a
        structured array of n (total_us, quantization_us, list_scan_us) rows."""
        kx, px = _ptr(x, np.float32, None, "x")
        n = kx.shape[0]
        D = _empty_like_input(kx, (n, k), np.float32)
        I = _empty_like_input(kx, (n, k), np.int64)
        kd, pd = _ptr(D, np.float32, (n, k), "D")
        ki, pi = _ptr(I, np.int64, (n, k), "I")
        self._order_with_torch(kx, kd, ki)
        st = (QueryLatencyStats * max(n, 1))()
        self._check(self._lib.b200_IndexIVFFlat_search_stats(
            self._h, n, px, k, pd, pi, params._h if params is not None else None, st))
        arr = np.frombuffer(st, dtype=[("total_us", "f8"), ("quantization_us", "f8"),
                                       ("list_scan_us", "f8")], count=n).copy()
        return kd, ki, arr

    def _range_result(self, call, n):
        rsr = _vp()
        self._check(self._lib.b200_RangeSearchResult_new(C.byref(rsr), n))
        try:
            self._check(call(rsr))
            lims_p = C.POINTER(C.c_size_t)()
            lab_p = C.POINTER(C.c_int64)()
            dis_p = C.POINTER(C.c_float)()
            self._lib.b200_RangeSearchResult_lims(rsr, C.byref(lims_p))
            self._lib.b200_RangeSearchResult_labels(rsr, C.byref(lab_p), C.byref(dis_p))
            lims = np.ctypeslib.as_array(lims_p, shape=(n + 1,)).astype(np.uint64).copy()
            tot = int(lims[-1])
            if tot:
                I = np.ctypeslib.as_array(lab_p, shape=(tot,)).copy()
                D = np.ctypeslib.as_array(dis_p, shape=(tot,)).copy()
            else:
                I, D = np.empty((0,), np.int64), np.empty((0,), np.float32)
            return lims, D, I
        finally:
            self._lib.b200_RangeSearchResult_free(rsr)

    def range_search(self, x, radius, params=None):
        """faiss Index.range_search -> (lims, D, I) (faiss/python/class_wrappers.py:370-401):
        L2 keeps dis < radius, inner product dis > radius."""
        kx, px = _ptr(x, np.float32, None, "x")
        n = kx.shape[0]
        self._order_with_torch(kx)
        return self._range_result(lambda r: self._lib.b200_IndexIVFFlat_range_search(
            self._h, n, px, float(radius), r, params._h if params is not None else None), n)

    def range_search_preassigned(self, x, radius, assign, store_pairs=False, params=None):
        kx, px = _ptr(x, np.float32, None, "x")
        n = kx.shape[0]
        nprobe = min(params.nprobe if params is not None else self.nprobe, self.nlist)
        ka, pa = _ptr(assign, np.int64, (n, nprobe), "assign")
        self._order_with_torch(kx, ka)
        return self._range_result(lambda r: self._lib.b200_IndexIVFFlat_range_search_preassigned(
            self._h, n, px, float(radius), pa, None, r, int(bool(store_pairs)),
            params._h if params is not None else None), n)

    def quantize(self, x, nprobe, Dq=None, Iq=None):
        kx, px = _ptr(x, np.float32, None, "x")
        n = kx.shape[0]
        if Dq is None:
            Dq = _empty_like_input(kx, (n, nprobe), np.float32)
        if Iq is None:
            Iq = _empty_like_input(kx, (n, nprobe), np.int64)
        kd, pd = _ptr(Dq, np.float32, (n, nprobe), "Dq")
        ki, pi = _ptr(Iq, np.int64, (n, nprobe), "Iq")
        self._order_with_torch(kx, kd, ki)
        self._check(self._lib.b200_IndexIVFFlat_quantize(self._h, n, px, nprobe, pd, pi))
        self._settle(kx, kd, ki)
        return kd, ki

    def quantize_noalloc(self, x, nprobe, Dq, Iq):
        """quantize() without any device allocation (b200ivf_c.h): CUDA tensors only, the caller
        supplies the outputs; n <= 4096, nprobe <= 64. Enqueued on the handle's stream."""
        kx, px = _ptr(x, np.float32, None, "x")
        n = kx.shape[0]
        kd, pd = _ptr(Dq, np.float32, (n, nprobe), "Dq")
        ki, pi = _ptr(Iq, np.int64, (n, nprobe), "Iq")
        self._order_with_torch(kx, kd, ki)
        self._check(self._lib.b200_IndexIVFFlat_quantize_noalloc(self._h, n, px, nprobe, pd, pi))
        self._settle(kx, kd, ki)
        return kd, ki

    # ---- stats / plumbing ----------------------------------------------------------------------
    def get_stats(self):
        s = IndexIVFStats()
        self._lib.b200_IndexIVFFlat_get_stats(self._h, C.byref(s))
        return s

    def reset_stats(self):
        self._lib.b200_IndexIVFFlat_reset_stats(self._h)

    def last_timings(self):
        ms = (C.c_double * 5)()
        self._lib.b200_IndexIVFFlat_last_timings(self._h, ms)
        return dict(zip(("coarse", "plan", "scan", "select", "total"), list(ms)))

    def stage_time_sums(self):
        """(calls, {stage: ms summed over the search calls since reset_stats})"""
        ms = (C.c_double * 5)()
        calls = self._lib.b200_IndexIVFFlat_stage_time_sums(self._h, ms)
        return int(calls), dict(zip(("coarse", "plan", "scan", "select", "total"), list(ms)))

    def exact_redo_count(self):
        """queries redone on the exact kernel so far (tc_split = 1, see b200ivf_c.h)"""
        return int(self._lib.b200_IndexIVFFlat_exact_redo_count(self._h))

    def set_option(self, name, value):
        """kernel-selection knobs (see b200ivf_c.h); results do not depend on them"""
        self._check(self._lib.b200_IndexIVFFlat_set_option(self._h, name.encode(), float(value)))

    def set_stream(self, cuda_stream_ptr):
        """Run on a caller-owned stream (e.g. ``torch.cuda.current_stream().cuda_stream``). torch's
        default stream has handle 0, which the C-ABI reads as "the handle's own stream": it is
        passed as cudaStreamLegacy (0x1) instead, so that the library really is stream-ordered
        with the caller's torch work. ``None`` selects the handle's own stream."""
        self._own_stream = cuda_stream_ptr is None
        if cuda_stream_ptr is None:
            cuda_stream_ptr = 0
        elif cuda_stream_ptr == 0:
            cuda_stream_ptr = 1  # cudaStreamLegacy
        self._check(self._lib.b200_IndexIVFFlat_set_stream(self._h, cuda_stream_ptr))


def read_index(fname, io_flags=0, device=0):
    """faiss.read_index for an IndexIVFFlat file written by the reference (or by write_index):
    the lists go straight into HBM on `device` (faiss/impl/index_read.cpp:782-787)."""
    lib = load_library()
    h = _vp()
    rc = lib.b200_read_index_fname(os.fsencode(fname), io_flags, device, C.byref(h))
    if rc != 0:
        raise B200Error(rc, lib.b200_get_last_error().decode())
    return IndexIVFFlat(0, 0, device=device, _handle=h)


def write_index(index, fname):
    """faiss.write_index in the reference's own format (faiss/impl/index_write.cpp:635-638)."""
    index._check(index._lib.b200_write_index_fname(index._h, os.fsencode(fname)))


def debug_trap_info():
    """None, or the note a stuck pipeline wait left before it trapped (see b200ivf_c.h)"""
    out = (C.c_ulonglong * 8)()
    if not load_library().b200_debug_trap_info(out):
        return None
    names = {1: "tma:empty", 2: "mma:d_empty", 3: "mma:a_rdy", 4: "split:full", 5: "split:a_empty",
             6: "epi:d_full"}
    return {"wait": names.get(int(out[0]) - 1, int(out[0]) - 1), "cta": int(out[1]),
            "thread": int(out[2]), "iteration": int(out[3]), "parity": int(out[4])}


def tensor_core_launches():
    """tensor-core scan launches so far (a subset of kernel_launches())"""
    return int(load_library().b200_tensor_core_launches())


def kernel_launches():
    """CUDA kernels launched by the library so far in this process."""
    return int(load_library().b200_kernel_launches())


def merge_knn_results(metric, all_D, all_I, device=0, stream=None):
    """merge_knn_results (faiss/utils/Heap.cpp:158-232) on the device: [nshard][n][k] -> [n][k]."""
    lib = load_library()
    kd, pd = _ptr(all_D, np.float32, None, "all_D")
    nshard, n, k = kd.shape
    ki, pi = _ptr(all_I, np.int64, (nshard, n, k), "all_I")
    D = _empty_like_input(kd, (n, k), np.float32)
    I = _empty_like_input(kd, (n, k), np.int64)
    ko, po = _ptr(D, np.float32, None, "D")
    kj, pj = _ptr(I, np.int64, None, "I")
    rc = lib.b200_merge_knn_results(metric, nshard, n, k, pd, pi, po, pj, device, stream)
    if rc != 0:
        raise B200Error(rc, lib.b200_get_last_error().decode())
    return ko, kj
