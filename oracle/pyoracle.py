"""oracle/pyoracle.py -- TEST INFRASTRUCTURE, not product code.

ctypes/numpy front-ends for the two CPU checkers:

* ``PortIndex``  -> oracle/liboracle_ivfflat.so, our plain-C restatement (ivfflat_oracle.c)
* ``RefIndex``   -> oracle/_ref/libref_ivfflat.so, the UNMODIFIED reference compiled from
                    /root/reference (ref_shim.cpp is only a C-ABI over its public C++ API)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module. The product package (hnsw_ivf_b200) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
METRIC_INNER_PRODUCT = 0
METRIC_L2 = 1

_f32p = C.POINTER(C.c_float)
_i64p = C.POINTER(C.c_int64)


def _f32(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a, a.ctypes.data_as(_f32p)


def _i64(a):
    a = np.ascontiguousarray(a, dtype=np.int64)
    return a, a.ctypes.data_as(_i64p)


# ------------------------------------------------------------------------------------------------
# port (plain-C restatement)
# ------------------------------------------------------------------------------------------------
class OrcStats(C.Structure):
    _fields_ = [("nq", C.c_size_t), ("nlist", C.c_size_t), ("ndis", C.c_size_t),
                ("nheap_updates", C.c_size_t), ("quantization_time", C.c_double),
                ("search_time", C.c_double)]


_port = None


def build_port():
    subprocess.check_call(["make", "-s", "-C", _HERE, "port"])


def port_lib():
    global _port
    if _port is None:
        path = os.path.join(_HERE, "liboracle_ivfflat.so")
        if not os.path.exists(path):
            build_port()
        L = C.CDLL(path)
        L.orc_index_new.restype = C.c_void_p
        L.orc_index_new.argtypes = [C.c_size_t, C.c_size_t, C.c_int]
        L.orc_index_free.argtypes = [C.c_void_p]
        L.orc_index_set_centroids.argtypes = [C.c_void_p, _f32p]
        L.orc_index_list_size.restype = C.c_size_t
        L.orc_index_list_size.argtypes = [C.c_void_p, C.c_size_t]
        L.orc_index_list_codes.restype = _f32p
        L.orc_index_list_codes.argtypes = [C.c_void_p, C.c_size_t]
        L.orc_index_list_ids.restype = _i64p
        L.orc_index_list_ids.argtypes = [C.c_void_p, C.c_size_t]
        L.orc_index_ntotal.restype = C.c_int64
        L.orc_index_ntotal.argtypes = [C.c_void_p]
        L.orc_index_add_core.argtypes = [C.c_void_p, C.c_int64, _f32p, _i64p, _i64p]
        L.orc_index_add_with_ids.argtypes = [C.c_void_p, C.c_int64, _f32p, _i64p]
        L.orc_index_quantize.argtypes = [C.c_void_p, C.c_int64, _f32p, C.c_size_t, _f32p, _i64p]
        L.orc_index_search_preassigned.restype = C.c_int
        L.orc_index_search_preassigned.argtypes = [C.c_void_p, C.c_int64, _f32p, C.c_int64,
                                                   C.c_size_t, _i64p, _f32p, _i64p,
                                                   C.POINTER(OrcStats)]
        L.orc_index_search_preassigned_ex.restype = C.c_int
        L.orc_index_search_preassigned_ex.argtypes = [C.c_void_p, C.c_int64, _f32p, C.c_int64,
                                                      C.c_size_t, _i64p, _f32p, _i64p,
                                                      C.POINTER(OrcStats), C.c_size_t, C.c_void_p,
                                                      C.c_int]
        L.orc_index_range_search_preassigned.restype = C.c_int
        L.orc_index_range_search_preassigned.argtypes = [
            C.c_void_p, C.c_int64, _f32p, C.c_float, C.c_size_t, _i64p, C.c_void_p, C.c_int,
            C.POINTER(C.c_size_t), C.POINTER(_i64p), C.POINTER(_f32p)]
        L.orc_free.argtypes = [C.c_void_p]
        L.orc_index_train.restype = C.c_int
        L.orc_index_train.argtypes = [C.c_void_p, C.c_int64, _f32p, C.c_int, C.c_int, C.c_int,
                                      C.POINTER(C.c_double)]
        L.orc_index_centroids.restype = _f32p
        L.orc_index_centroids.argtypes = [C.c_void_p]
        L.orc_sel_range.restype = C.c_void_p
        L.orc_sel_range.argtypes = [C.c_int64, C.c_int64]
        L.orc_sel_batch.restype = C.c_void_p
        L.orc_sel_batch.argtypes = [C.c_size_t, _i64p]
        L.orc_sel_bitmap.restype = C.c_void_p
        L.orc_sel_bitmap.argtypes = [C.c_size_t, C.c_void_p]
        L.orc_sel_combine.restype = C.c_void_p
        L.orc_sel_combine.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        L.orc_sel_free.argtypes = [C.c_void_p]
        L.orc_sel_is_member.restype = C.c_int
        L.orc_sel_is_member.argtypes = [C.c_void_p, C.c_int64]
        L.orc_index_search.restype = C.c_int
        L.orc_index_search.argtypes = [C.c_void_p, C.c_int64, _f32p, C.c_int64, C.c_size_t,
                                       C.c_int, _f32p, _i64p, C.POINTER(OrcStats)]
        L.orc_knn.argtypes = [C.c_int, _f32p, _f32p, C.c_size_t, C.c_size_t, C.c_size_t,
                              C.c_size_t, _f32p, _i64p]
        L.orc_merge_knn_results.argtypes = [C.c_int, C.c_size_t, C.c_size_t, C.c_int, _f32p,
                                            _i64p, _f32p, _i64p]
        L.orc_heap_heapify.argtypes = [C.c_int, C.c_size_t, _f32p, _i64p]
        L.orc_heap_addn.restype = C.c_size_t
        L.orc_heap_addn.argtypes = [C.c_int, C.c_size_t, _f32p, _i64p, _f32p, _i64p, C.c_size_t]
        L.orc_heap_reorder.restype = C.c_size_t
        L.orc_heap_reorder.argtypes = [C.c_int, C.c_size_t, _f32p, _i64p]
        for fn in (L.orc_fvec_L2sqr, L.orc_fvec_inner_product):
            fn.restype = C.c_float
            fn.argtypes = [_f32p, _f32p, C.c_size_t]
        _port = L
    return _port


# ---- selector specs -------------------------------------------------------------------------------
# One description of an IDSelector tree, understood by the port, the reference shim and the tests'
# GPU builder: ("range", imin, imax) | ("batch", ids) | ("bitmap", uint8 array) | ("not", s) |
# ("and", a, b) | ("or", a, b) | ("xor", a, b)
_COMBINE = {"not": 3, "and": 4, "or": 5, "xor": 6}


class _SelTree:
    """Owns the C objects of one selector tree (children before parents) and frees them."""

    def __init__(self, spec, leaf_range, leaf_batch, leaf_bitmap, combine, free):
        self._free = free
        self._nodes = []

        def build(sp):
            kind = sp[0]
            if kind == "range":
                h = leaf_range(int(sp[1]), int(sp[2]))
            elif kind == "batch":
                a = np.ascontiguousarray(sp[1], dtype=np.int64)
                h = leaf_batch(a.size, a.ctypes.data_as(_i64p))
            elif kind == "bitmap":
                a = np.ascontiguousarray(sp[1], dtype=np.uint8)
                h = leaf_bitmap(a.size, a.ctypes.data_as(C.c_void_p))
            else:
                lhs = build(sp[1])
                rhs = build(sp[2]) if kind != "not" else None
                h = combine(_COMBINE[kind], lhs, rhs)
            h = C.c_void_p(h)
            self._nodes.append(h)
            return h

        self.h = build(spec) if spec is not None else None

    def __del__(self):
        for h in reversed(getattr(self, "_nodes", [])):
            self._free(h)
        self._nodes = []


def port_sel(spec):
    L = port_lib()
    return _SelTree(spec, L.orc_sel_range, L.orc_sel_batch, L.orc_sel_bitmap, L.orc_sel_combine,
                    L.orc_sel_free)


def sel_is_member_py(spec, i):
    """pure-Python statement of the same predicates (faiss/impl/IDSelector.cpp:17-123)"""
    kind = spec[0]
    if kind == "range":
        return spec[1] <= i < spec[2]
    if kind == "batch":
        return int(i) in set(int(v) for v in spec[1])
    if kind == "bitmap":
        b = np.asarray(spec[1], np.uint8)
        u = int(i) & 0xFFFFFFFFFFFFFFFF
        return (u >> 3) < b.size and bool((int(b[u >> 3]) >> (u & 7)) & 1)
    if kind == "not":
        return not sel_is_member_py(spec[1], i)
    a, b = sel_is_member_py(spec[1], i), sel_is_member_py(spec[2], i)
    return (a and b) if kind == "and" else (a or b) if kind == "or" else (a != b)


class PortIndex:
    """Restatement of IndexIVFFlat + IndexFlat quantizer (oracle/ivfflat_oracle.c)."""

    def __init__(self, d, nlist, metric):
        self.L = port_lib()
        self.d, self.nlist, self.metric = int(d), int(nlist), int(metric)
        self.h = C.c_void_p(self.L.orc_index_new(self.d, self.nlist, self.metric))
        self.stats = OrcStats()

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_index_free(self.h)
            self.h = None

    def set_centroids(self, c):
        c, p = _f32(c)
        assert c.shape == (self.nlist, self.d)
        self.L.orc_index_set_centroids(self.h, p)

    @property
    def ntotal(self):
        return self.L.orc_index_ntotal(self.h)

    def add_with_ids(self, x, ids=None):
        x, px = _f32(x)
        pi = None
        if ids is not None:
            ids, pi = _i64(ids)
        self.L.orc_index_add_with_ids(self.h, x.shape[0], px, pi)

    def add_core(self, x, ids, assign):
        x, px = _f32(x)
        pi = None
        if ids is not None:
            ids, pi = _i64(ids)
        assign, pa = _i64(assign)
        self.L.orc_index_add_core(self.h, x.shape[0], px, pi, pa)

    def list_size(self, l):
        return self.L.orc_index_list_size(self.h, l)

    def get_list(self, l):
        n = self.list_size(l)
        if n == 0:
            return np.zeros((0, self.d), np.float32), np.zeros((0,), np.int64)
        codes = np.ctypeslib.as_array(self.L.orc_index_list_codes(self.h, l), (n, self.d)).copy()
        ids = np.ctypeslib.as_array(self.L.orc_index_list_ids(self.h, l), (n,)).copy()
        return codes, ids

    def quantize(self, x, nprobe):
        x, px = _f32(x)
        n = x.shape[0]
        nprobe = min(nprobe, self.nlist)
        Dq = np.empty((n, nprobe), np.float32)
        Iq = np.empty((n, nprobe), np.int64)
        self.L.orc_index_quantize(self.h, n, px, nprobe, Dq.ctypes.data_as(_f32p),
                                  Iq.ctypes.data_as(_i64p))
        return Dq, Iq

    def search_preassigned(self, x, k, keys):
        x, px = _f32(x)
        keys, pk = _i64(keys)
        n, nprobe = keys.shape
        D = np.empty((n, k), np.float32)
        I = np.empty((n, k), np.int64)
        rc = self.L.orc_index_search_preassigned(self.h, n, px, k, nprobe, pk,
                                                 D.ctypes.data_as(_f32p),
                                                 I.ctypes.data_as(_i64p), C.byref(self.stats))
        if rc:
            raise RuntimeError("oracle search_preassigned failed rc=%d" % rc)
        return D, I

    def search_preassigned_ex(self, x, k, keys, max_codes=0, sel=None, store_pairs=False):
        """search_preassigned with the SearchParametersIVF arms; sel is a selector spec"""
        x, px = _f32(x)
        keys, pk = _i64(keys)
        n, nprobe = keys.shape
        D = np.empty((n, k), np.float32)
        I = np.empty((n, k), np.int64)
        tree = port_sel(sel)
        rc = self.L.orc_index_search_preassigned_ex(
            self.h, n, px, k, nprobe, pk, D.ctypes.data_as(_f32p), I.ctypes.data_as(_i64p),
            C.byref(self.stats), int(max_codes), tree.h, int(bool(store_pairs)))
        if rc:
            raise RuntimeError("oracle search_preassigned_ex failed rc=%d" % rc)
        return D, I

    def range_search_preassigned(self, x, radius, keys, sel=None, store_pairs=False):
        x, px = _f32(x)
        keys, pk = _i64(keys)
        n, nprobe = keys.shape
        lims = (C.c_size_t * (n + 1))()
        lab, dis = _i64p(), _f32p()
        tree = port_sel(sel)
        rc = self.L.orc_index_range_search_preassigned(
            self.h, n, px, float(radius), nprobe, pk, tree.h, int(bool(store_pairs)), lims,
            C.byref(lab), C.byref(dis))
        try:
            if rc:
                raise RuntimeError("oracle range_search_preassigned failed rc=%d" % rc)
            lims = np.array(list(lims), dtype=np.uint64)
            tot = int(lims[-1])
            I = np.ctypeslib.as_array(lab, (tot,)).copy() if tot else np.empty((0,), np.int64)
            D = np.ctypeslib.as_array(dis, (tot,)).copy() if tot else np.empty((0,), np.float32)
        finally:
            self.L.orc_free(lab)
            self.L.orc_free(dis)
        return lims, D, I

    def train(self, x, niter=10, seed=1234, max_points_per_centroid=256):
        """k-means as IndexIVF::train runs it; returns the last iteration's objective"""
        x, px = _f32(x)
        obj = C.c_double(0)
        rc = self.L.orc_index_train(self.h, x.shape[0], px, niter, seed, max_points_per_centroid,
                                    C.byref(obj))
        if rc:
            raise RuntimeError("oracle train failed rc=%d" % rc)
        return obj.value

    def get_centroids(self):
        return np.ctypeslib.as_array(self.L.orc_index_centroids(self.h),
                                     (self.nlist, self.d)).copy()

    def search(self, x, k, nprobe, nt=1):
        x, px = _f32(x)
        n = x.shape[0]
        D = np.empty((n, k), np.float32)
        I = np.empty((n, k), np.int64)
        rc = self.L.orc_index_search(self.h, n, px, k, nprobe, nt, D.ctypes.data_as(_f32p),
                                     I.ctypes.data_as(_i64p), C.byref(self.stats))
        if rc:
            raise RuntimeError("oracle search failed rc=%d" % rc)
        return D, I


def port_knn(metric, x, y, k):
    L = port_lib()
    x, px = _f32(x)
    y, py = _f32(y)
    D = np.empty((x.shape[0], k), np.float32)
    I = np.empty((x.shape[0], k), np.int64)
    L.orc_knn(metric, px, py, x.shape[1], x.shape[0], y.shape[0], k, D.ctypes.data_as(_f32p),
              I.ctypes.data_as(_i64p))
    return D, I


def port_merge_knn_results(metric, all_D, all_I):
    L = port_lib()
    all_D, pd = _f32(all_D)
    all_I, pi = _i64(all_I)
    nshard, n, k = all_D.shape
    D = np.empty((n, k), np.float32)
    I = np.empty((n, k), np.int64)
    L.orc_merge_knn_results(metric, n, k, nshard, pd, pi, D.ctypes.data_as(_f32p),
                            I.ctypes.data_as(_i64p))
    return D, I


def port_heap_select(metric, k, vals, ids):
    """heapify + addn + reorder: the reference's top-k semantics (strict compare, cmp2 order)."""
    L = port_lib()
    vals, pv = _f32(vals)
    ids, pi = _i64(ids)
    ov = np.empty(k, np.float32)
    oi = np.empty(k, np.int64)
    L.orc_heap_heapify(metric, k, ov.ctypes.data_as(_f32p), oi.ctypes.data_as(_i64p))
    L.orc_heap_addn(metric, k, ov.ctypes.data_as(_f32p), oi.ctypes.data_as(_i64p), pv, pi,
                    vals.shape[0])
    L.orc_heap_reorder(metric, k, ov.ctypes.data_as(_f32p), oi.ctypes.data_as(_i64p))
    return ov, oi


# ------------------------------------------------------------------------------------------------
# compiled reference (oracle/_ref)
# ------------------------------------------------------------------------------------------------
_ref = None
_REF_DIR = os.path.join(_HERE, "_ref")


def ref_available():
    return os.path.exists(os.path.join(_REF_DIR, "libref_ivfflat.so"))


def build_ref(reference_root="/root/reference"):
    """Compile the reference from its own sources (only possible where they are mounted)."""
    if not os.path.isdir(reference_root):
        raise RuntimeError("reference sources not present; using prebuilt oracle/_ref if any")
    subprocess.check_call(["make", "-s", "-j8", "-C", _HERE, "ref", "REF=" + reference_root])


def ref_lib():
    global _ref
    if _ref is None:
        if not ref_available():
            raise RuntimeError("oracle/_ref/libref_ivfflat.so missing (run oracle/Makefile 'ref')")
        # OpenBLAS inside faiss is called from faiss' own OpenMP slices: one BLAS thread each
        os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")
        for dep in ("libquadmath-2284e583.so.0.0.0", "libgfortran-83c28eba.so.5.0.0"):
            p = os.path.join(_REF_DIR, dep)
            if os.path.exists(p):
                C.CDLL(p, mode=C.RTLD_GLOBAL)
        for f in sorted(os.listdir(_REF_DIR)):
            if f.startswith("libopenblas"):
                C.CDLL(os.path.join(_REF_DIR, f), mode=C.RTLD_GLOBAL)
        L = C.CDLL(os.path.join(_REF_DIR, "libref_ivfflat.so"))
        L.ref_last_error.restype = C.c_char_p
        L.ref_getmillisecs.restype = C.c_double
        L.ref_ivfflat_new.restype = C.c_void_p
        L.ref_ivfflat_new.argtypes = [C.c_int64, C.c_int64, C.c_int]
        L.ref_ivfflat_free.argtypes = [C.c_void_p]
        L.ref_ivfflat_set_centroids.argtypes = [C.c_void_p, _f32p]
        L.ref_ivfflat_train.argtypes = [C.c_void_p, C.c_int64, _f32p]
        L.ref_ivfflat_get_centroids.argtypes = [C.c_void_p, _f32p]
        L.ref_ivfflat_add_with_ids.argtypes = [C.c_void_p, C.c_int64, _f32p, _i64p]
        L.ref_ivfflat_add_core.argtypes = [C.c_void_p, C.c_int64, _f32p, _i64p, _i64p]
        L.ref_ivfflat_append_list.argtypes = [C.c_void_p, C.c_int64, C.c_int64, _i64p, _f32p]
        L.ref_ivfflat_ntotal.restype = C.c_int64
        L.ref_read_index.restype = C.c_void_p
        L.ref_read_index.argtypes = [C.c_char_p, _i64p]
        L.ref_ivfflat_write_index.argtypes = [C.c_void_p, C.c_char_p]
        L.ref_ivfflat_set_nprobe.argtypes = [C.c_void_p, C.c_int64]
        L.ref_ivfflat_ntotal.argtypes = [C.c_void_p]
        L.ref_ivfflat_list_size.restype = C.c_int64
        L.ref_ivfflat_list_size.argtypes = [C.c_void_p, C.c_int64]
        L.ref_ivfflat_get_list.argtypes = [C.c_void_p, C.c_int64, _f32p, _i64p]
        L.ref_ivfflat_quantize.argtypes = [C.c_void_p, C.c_int64, _f32p, C.c_int64, _f32p, _i64p]
        L.ref_ivfflat_search.argtypes = [C.c_void_p, C.c_int64, _f32p, C.c_int64, C.c_int64,
                                         _f32p, _i64p]
        L.ref_ivfflat_search_preassigned.argtypes = [C.c_void_p, C.c_int64, _f32p, C.c_int64,
                                                     C.c_int64, _i64p, _f32p, _f32p, _i64p]
        L.ref_ivfflat_set_parallel_mode.argtypes = [C.c_void_p, C.c_int]
        L.ref_sel_range_new.restype = C.c_void_p
        L.ref_sel_range_new.argtypes = [C.c_int64, C.c_int64]
        L.ref_sel_batch_new.restype = C.c_void_p
        L.ref_sel_batch_new.argtypes = [C.c_int64, _i64p]
        L.ref_sel_bitmap_new.restype = C.c_void_p
        L.ref_sel_bitmap_new.argtypes = [C.c_int64, C.c_void_p]
        L.ref_sel_combine_new.restype = C.c_void_p
        L.ref_sel_combine_new.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        L.ref_sel_is_member.argtypes = [C.c_void_p, C.c_int64]
        L.ref_sel_free.argtypes = [C.c_void_p]
        L.ref_ivfflat_search_params.argtypes = [C.c_void_p, C.c_int64, _f32p, C.c_int64, C.c_int64,
                                                C.c_int64, C.c_void_p, _f32p, _i64p]
        L.ref_ivfflat_search_preassigned_ex.argtypes = [
            C.c_void_p, C.c_int64, _f32p, C.c_int64, C.c_int64, _i64p, _f32p, _f32p, _i64p,
            C.c_int, C.c_int64, C.c_void_p]
        L.ref_ivfflat_search_stats.argtypes = [C.c_void_p, C.c_int64, _f32p, C.c_int64, C.c_int64,
                                               _f32p, _i64p, C.POINTER(C.c_double)]
        L.ref_ivfflat_range_search.restype = C.c_void_p
        L.ref_ivfflat_range_search.argtypes = [C.c_void_p, C.c_int64, _f32p, C.c_float, C.c_int64,
                                               _i64p, C.c_int, C.c_void_p]
        L.ref_range_lims.restype = C.POINTER(C.c_size_t)
        L.ref_range_lims.argtypes = [C.c_void_p]
        L.ref_range_labels.restype = _i64p
        L.ref_range_labels.argtypes = [C.c_void_p]
        L.ref_range_distances.restype = _f32p
        L.ref_range_distances.argtypes = [C.c_void_p]
        L.ref_range_free.argtypes = [C.c_void_p]
        L.ref_ivfflat_set_cp.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.ref_stats_get.argtypes = [C.POINTER(C.c_double)]
        L.ref_knn.argtypes = [C.c_int, _f32p, _f32p, C.c_int64, C.c_int64, C.c_int64, C.c_int64,
                              _f32p, _i64p]
        L.ref_merge_knn_results.argtypes = [C.c_int, C.c_int64, C.c_int64, C.c_int, _f32p, _i64p,
                                            _f32p, _i64p]
        L.ref_heap_select.argtypes = [C.c_int, C.c_int64, C.c_int64, _f32p, _i64p, _f32p, _i64p]
        for fn in (L.ref_fvec_L2sqr, L.ref_fvec_inner_product):
            fn.restype = C.c_float
            fn.argtypes = [_f32p, _f32p, C.c_int64]
        _ref = L
    return _ref


def ref_sel(spec):
    L = ref_lib()
    return _SelTree(spec, L.ref_sel_range_new, L.ref_sel_batch_new, L.ref_sel_bitmap_new,
                    L.ref_sel_combine_new, L.ref_sel_free)


def _chk(L, rc):
    if rc:
        raise RuntimeError("reference error %d: %s" % (rc, L.ref_last_error().decode()))


class RefIndex:
    """The reference's own IndexIVFFlat + IndexFlat quantizer (faiss/IndexIVFFlat.h:26)."""

    def __init__(self, d, nlist, metric):
        self.L = ref_lib()
        self.d, self.nlist, self.metric = int(d), int(nlist), int(metric)
        self.h = C.c_void_p(self.L.ref_ivfflat_new(self.d, self.nlist, self.metric))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_ivfflat_free(self.h)
            self.h = None

    def set_centroids(self, c):
        c, p = _f32(c)
        assert c.shape == (self.nlist, self.d)
        _chk(self.L, self.L.ref_ivfflat_set_centroids(self.h, p))

    def train(self, x):
        x, p = _f32(x)
        _chk(self.L, self.L.ref_ivfflat_train(self.h, x.shape[0], p))

    def get_centroids(self):
        c = np.empty((self.nlist, self.d), np.float32)
        _chk(self.L, self.L.ref_ivfflat_get_centroids(self.h, c.ctypes.data_as(_f32p)))
        return c

    @property
    def ntotal(self):
        return self.L.ref_ivfflat_ntotal(self.h)

    def add_with_ids(self, x, ids=None):
        x, px = _f32(x)
        pi = None
        if ids is not None:
            ids, pi = _i64(ids)
        _chk(self.L, self.L.ref_ivfflat_add_with_ids(self.h, x.shape[0], px, pi))

    def add_core(self, x, ids, assign):
        x, px = _f32(x)
        pi = None
        if ids is not None:
            ids, pi = _i64(ids)
        assign, pa = _i64(assign)
        _chk(self.L, self.L.ref_ivfflat_add_core(self.h, x.shape[0], px, pi, pa))

    def append_list(self, l, ids, codes):
        codes, pc = _f32(codes)
        ids, pi = _i64(ids)
        _chk(self.L, self.L.ref_ivfflat_append_list(self.h, l, ids.shape[0], pi, pc))

    def list_size(self, l):
        return self.L.ref_ivfflat_list_size(self.h, l)

    def get_list(self, l):
        n = self.list_size(l)
        codes = np.empty((n, self.d), np.float32)
        ids = np.empty((n,), np.int64)
        if n:
            _chk(self.L, self.L.ref_ivfflat_get_list(self.h, l, codes.ctypes.data_as(_f32p),
                                                     ids.ctypes.data_as(_i64p)))
        return codes, ids

    def quantize(self, x, nprobe):
        x, px = _f32(x)
        n = x.shape[0]
        Dq = np.empty((n, nprobe), np.float32)
        Iq = np.empty((n, nprobe), np.int64)
        _chk(self.L, self.L.ref_ivfflat_quantize(self.h, n, px, nprobe,
                                                 Dq.ctypes.data_as(_f32p),
                                                 Iq.ctypes.data_as(_i64p)))
        return Dq, Iq

    def search(self, x, k, nprobe):
        x, px = _f32(x)
        n = x.shape[0]
        D = np.empty((n, k), np.float32)
        I = np.empty((n, k), np.int64)
        _chk(self.L, self.L.ref_ivfflat_search(self.h, n, px, k, nprobe,
                                               D.ctypes.data_as(_f32p), I.ctypes.data_as(_i64p)))
        return D, I

    def search_preassigned(self, x, k, keys, cdis=None):
        x, px = _f32(x)
        keys, pk = _i64(keys)
        n, nprobe = keys.shape
        if cdis is None:
            cdis = np.zeros((n, nprobe), np.float32)
        cdis, pc = _f32(cdis)
        D = np.empty((n, k), np.float32)
        I = np.empty((n, k), np.int64)
        _chk(self.L, self.L.ref_ivfflat_search_preassigned(self.h, n, px, k, nprobe, pk, pc,
                                                           D.ctypes.data_as(_f32p),
                                                           I.ctypes.data_as(_i64p)))
        return D, I

    def search_params(self, x, k, nprobe, max_codes=0, sel=None):
        """IndexIVF::search with SearchParametersIVF; sel is a selector spec"""
        x, px = _f32(x)
        n = x.shape[0]
        D = np.empty((n, k), np.float32)
        I = np.empty((n, k), np.int64)
        tree = ref_sel(sel)
        _chk(self.L, self.L.ref_ivfflat_search_params(self.h, n, px, k, nprobe, int(max_codes),
                                                      tree.h, D.ctypes.data_as(_f32p),
                                                      I.ctypes.data_as(_i64p)))
        return D, I

    def search_preassigned_ex(self, x, k, keys, max_codes=0, sel=None, store_pairs=False):
        x, px = _f32(x)
        keys, pk = _i64(keys)
        n, nprobe = keys.shape
        cdis, pc = _f32(np.zeros((n, nprobe), np.float32))
        D = np.empty((n, k), np.float32)
        I = np.empty((n, k), np.int64)
        tree = ref_sel(sel)
        _chk(self.L, self.L.ref_ivfflat_search_preassigned_ex(
            self.h, n, px, k, nprobe, pk, pc, D.ctypes.data_as(_f32p), I.ctypes.data_as(_i64p),
            int(bool(store_pairs)), int(max_codes), tree.h))
        return D, I

    def search_stats(self, x, k, nprobe):
        """the fork's IndexIVF::search_stats -> (D, I, [n][3] total/quantization/list_scan us)"""
        x, px = _f32(x)
        n = x.shape[0]
        D = np.empty((n, k), np.float32)
        I = np.empty((n, k), np.int64)
        st = np.zeros((n, 3), np.float64)
        _chk(self.L, self.L.ref_ivfflat_search_stats(
            self.h, n, px, k, nprobe, D.ctypes.data_as(_f32p), I.ctypes.data_as(_i64p),
            st.ctypes.data_as(C.POINTER(C.c_double))))
        return D, I, st

    def range_search(self, x, radius, nprobe, keys=None, sel=None, store_pairs=False):
        """IndexIVF::range_search (keys None) / range_search_preassigned -> (lims, D, I)"""
        x, px = _f32(x)
        n = x.shape[0]
        pk = None
        if keys is not None:
            keys, pk = _i64(keys)
        tree = ref_sel(sel)
        res = self.L.ref_ivfflat_range_search(self.h, n, px, float(radius), nprobe, pk,
                                              int(bool(store_pairs)), tree.h)
        if not res:
            raise RuntimeError("reference range_search: %s" % self.L.ref_last_error().decode())
        res = C.c_void_p(res)
        try:
            lims = np.ctypeslib.as_array(self.L.ref_range_lims(res), (n + 1,)).astype(np.uint64)
            tot = int(lims[-1])
            if tot:
                I = np.ctypeslib.as_array(self.L.ref_range_labels(res), (tot,)).copy()
                D = np.ctypeslib.as_array(self.L.ref_range_distances(res), (tot,)).copy()
            else:
                I, D = np.empty((0,), np.int64), np.empty((0,), np.float32)
        finally:
            self.L.ref_range_free(res)
        return lims, D, I

    def set_cp(self, niter=10, seed=1234, max_points_per_centroid=256):
        self.L.ref_ivfflat_set_cp(self.h, niter, seed, max_points_per_centroid)

    def set_parallel_mode(self, pm):
        self.L.ref_ivfflat_set_parallel_mode(self.h, pm)

    def set_nprobe(self, nprobe):
        """the nprobe FIELD (what write_index stores); search() here takes nprobe as an argument"""
        self.L.ref_ivfflat_set_nprobe(self.h, int(nprobe))

    def write_index(self, fname):
        """faiss::write_index (faiss/impl/index_write.cpp:391,635-638)"""
        _chk(self.L, self.L.ref_ivfflat_write_index(self.h, os.fsencode(fname)))

    @classmethod
    def read_index(cls, fname):
        """faiss::read_index (faiss/impl/index_read.cpp:593,782-787) of an IndexIVFFlat file"""
        L = ref_lib()
        out = np.zeros(5, np.int64)
        h = L.ref_read_index(os.fsencode(fname), out.ctypes.data_as(_i64p))
        if not h:
            raise RuntimeError("reference read_index failed: %s" % L.ref_last_error().decode())
        r = cls.__new__(cls)
        r.L = L
        r.d, r.nlist, r.metric = int(out[0]), int(out[1]), int(out[2])
        r.nprobe = int(out[3])
        r.h = C.c_void_p(h)
        return r


def ref_stats(reset=False):
    L = ref_lib()
    out = (C.c_double * 6)()
    L.ref_stats_get(out)
    if reset:
        L.ref_stats_reset()
    keys = ("nq", "nlist", "ndis", "nheap_updates", "quantization_time", "search_time")
    return dict(zip(keys, list(out)))


def ref_set_threads(nt):
    ref_lib().ref_set_omp_threads(int(nt))


def ref_max_threads():
    return ref_lib().ref_omp_max_threads()


def ref_knn(metric, x, y, k):
    L = ref_lib()
    x, px = _f32(x)
    y, py = _f32(y)
    D = np.empty((x.shape[0], k), np.float32)
    I = np.empty((x.shape[0], k), np.int64)
    _chk(L, L.ref_knn(metric, px, py, x.shape[1], x.shape[0], y.shape[0], k,
                      D.ctypes.data_as(_f32p), I.ctypes.data_as(_i64p)))
    return D, I


def ref_merge_knn_results(metric, all_D, all_I):
    L = ref_lib()
    all_D, pd = _f32(all_D)
    all_I, pi = _i64(all_I)
    nshard, n, k = all_D.shape
    D = np.empty((n, k), np.float32)
    I = np.empty((n, k), np.int64)
    _chk(L, L.ref_merge_knn_results(metric, n, k, nshard, pd, pi, D.ctypes.data_as(_f32p),
                                    I.ctypes.data_as(_i64p)))
    return D, I


def ref_heap_select(metric, k, vals, ids):
    L = ref_lib()
    vals, pv = _f32(vals)
    ids, pi = _i64(ids)
    ov = np.empty(k, np.float32)
    oi = np.empty(k, np.int64)
    L.ref_heap_select(metric, k, vals.shape[0], pv, pi, ov.ctypes.data_as(_f32p),
                      oi.ctypes.data_as(_i64p))
    return ov, oi


# ---- the reference's index file format, parsed with numpy (faiss/impl/index_write.cpp) -----------
def _fourcc(s):
    """faiss/impl/io.cpp:237-241"""
    b = s.encode()
    return b[0] | b[1] << 8 | b[2] << 16 | b[3] << 24


def parse_index_file(path):
    """Plain restatement of read_index for one index type: "IwFl" (faiss/impl/index_read.cpp:
    782-787) = index header (:62-74) + nlist + nprobe + IndexFlat quantizer ("IxF2"/"IxFI", :593-611)
    + direct map (:442-456) + "ilar" ArrayInvertedLists (:155-212). Returns a dict with the header
    fields, the centroids and the lists as (codes [n][d] fp32, ids [n] int64)."""
    buf = np.fromfile(path, dtype=np.uint8)
    pos = [0]

    def take(dtype, n=1):
        dt = np.dtype(dtype)
        a = np.frombuffer(buf, dtype=dt, count=n, offset=pos[0])
        pos[0] += dt.itemsize * n
        return a

    def header():
        d = int(take("<i4")[0])
        ntotal = int(take("<i8")[0])
        take("<i8", 2)  # two dummies
        trained = bool(take("u1")[0])
        metric = int(take("<i4")[0])
        assert metric in (0, 1), "metric_arg follows for other metrics"
        return d, ntotal, trained, metric

    h = int(take("<u4")[0])
    assert h == _fourcc("IwFl"), "not an IndexIVFFlat file"
    d, ntotal, trained, metric = header()
    nlist, nprobe = int(take("<u8")[0]), int(take("<u8")[0])
    qh = int(take("<u4")[0])
    assert qh in (_fourcc("IxF2"), _fourcc("IxFI"), _fourcc("IxFl"))
    qd, qn, qtrained, qmetric = header()
    nfl = int(take("<u8")[0])
    centroids = take("<f4", nfl).reshape(-1, qd).copy()
    dm_type = int(take("u1")[0])
    take("<i8", int(take("<u8")[0]))
    if dm_type == 2:
        take("<i8", 2 * int(take("<u8")[0]))
    ih = int(take("<u4")[0])
    lists = None
    if ih == _fourcc("ilar"):
        il_nlist, code_size = int(take("<u8")[0]), int(take("<u8")[0])
        assert il_nlist == nlist and code_size == 4 * d
        lt = int(take("<u4")[0])
        tab = take("<u8", int(take("<u8")[0]))
        sizes = np.zeros(nlist, np.int64)
        if lt == _fourcc("full"):
            sizes[:] = tab
        else:
            assert lt == _fourcc("sprs")
            sizes[tab[0::2].astype(np.int64)] = tab[1::2]
        lists = []
        for l in range(nlist):
            n = int(sizes[l])
            codes = take("<f4", n * d).reshape(n, d).copy()
            ids = take("<i8", n).copy()
            lists.append((codes, ids))
    else:
        assert ih == _fourcc("il00")
    assert pos[0] == buf.size, "trailing bytes"
    return dict(d=d, ntotal=ntotal, is_trained=trained, metric=metric, nlist=nlist, nprobe=nprobe,
                quantizer=dict(fourcc=qh, d=qd, ntotal=qn, is_trained=qtrained, metric=qmetric),
                centroids=centroids, direct_map_type=dm_type, lists=lists)
