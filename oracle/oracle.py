"""ctypes binding of the CPU oracle (oracle/liborc.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs.  The product package (synapseml_b200/) never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    so = os.path.join(_HERE, "liborc.so")
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".h", ".cpp"))]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liborc.so")
        if not os.path.exists(so):
            build()
        _LIB = C.CDLL(so)
        _LIB.ORC_GetLastError.restype = C.c_char_p
        _LIB.ORC_SoftExp.restype = C.c_double
        _LIB.ORC_SoftExp.argtypes = [C.c_double]
        _LIB.ORC_SoftLog2.restype = C.c_double
        _LIB.ORC_SoftLog2.argtypes = [C.c_double]
        _LIB.ORC_HashUniform.restype = C.c_float
        _LIB.ORC_HashUniform.argtypes = [C.c_uint32] * 4
    return _LIB


def _check(rc):
    if rc != 0:
        raise RuntimeError("oracle call failed: " + lib().ORC_GetLastError().decode())


def _ptr(a, t):
    return a.ctypes.data_as(C.POINTER(t)) if a is not None else None


def find_bin(values, total_sample_cnt, max_bin=255, min_data_in_bin=3, min_split_data=0, pre_filter=True,
             use_missing=True, zero_as_missing=False):
    """BinMapper::FindBin on the non-zero sampled values of one feature."""
    v = np.ascontiguousarray(values, dtype=np.float64)
    bounds = np.zeros(max(max_bin, 2) + 2, dtype=np.float64)
    nb, mt, db, mf, tr = (C.c_int() for _ in range(5))
    _check(lib().ORC_FindBin(_ptr(v, C.c_double), C.c_int(len(v)), C.c_int(total_sample_cnt), C.c_int(max_bin),
                             C.c_int(min_data_in_bin), C.c_int(min_split_data), C.c_int(pre_filter), C.c_int(use_missing),
                             C.c_int(zero_as_missing), _ptr(bounds, C.c_double), C.byref(nb), C.byref(mt), C.byref(db),
                             C.byref(mf), C.byref(tr)))
    return dict(bounds=bounds[:nb.value].copy(), num_bin=nb.value, missing_type=mt.value, default_bin=db.value,
                most_freq_bin=mf.value, is_trivial=bool(tr.value))


def sampled_columns(sample):
    """Column-major non-zero compaction of a dense sample (mirror of SampledData.pushRow, dataset/SampledData.scala:43-77)."""
    sample = np.asarray(sample, dtype=np.float64)
    n, f = sample.shape
    cols, idxs, counts = [], [], np.zeros(f, dtype=np.int32)
    for j in range(f):
        c = sample[:, j]
        nz = np.flatnonzero(c != 0.0)      # NaN != 0.0 is True: NaNs are kept
        cols.append(np.ascontiguousarray(c[nz]))
        idxs.append(nz.astype(np.int32))
        counts[j] = len(nz)
    return cols, idxs, counts


class Dataset:
    def __init__(self, handle):
        self.h = handle

    @staticmethod
    def from_sample(sample, num_local_row, num_dist_row, params):
        cols, idxs, counts = sampled_columns(sample)
        f = len(cols)
        dpp = (C.POINTER(C.c_double) * f)(*[_ptr(c, C.c_double) for c in cols])
        ipp = (C.POINTER(C.c_int) * f)(*[_ptr(i, C.c_int) for i in idxs])
        out = C.c_void_p()
        _check(lib().ORC_DatasetCreateFromSampledColumn(dpp, ipp, C.c_int32(f), _ptr(counts, C.c_int), C.c_int32(sample.shape[0]),
                                                        C.c_int32(num_local_row), C.c_int64(num_dist_row), params.encode(),
                                                        C.byref(out)))
        return Dataset(out)

    def create_by_reference(self, num_rows):
        out = C.c_void_p()
        _check(lib().ORC_DatasetCreateByReference(self.h, C.c_int64(num_rows), C.byref(out)))
        return Dataset(out)

    def init_streaming(self, has_weights=False, has_init_scores=False, has_queries=False):
        _check(lib().ORC_DatasetInitStreaming(self.h, C.c_int32(has_weights), C.c_int32(has_init_scores), C.c_int32(has_queries),
                                              C.c_int32(1), C.c_int32(1), C.c_int32(1)))

    def push_rows(self, data, start_row, label, weight=None, init_score=None, query=None):
        data = np.ascontiguousarray(data, dtype=np.float64)
        label = np.ascontiguousarray(label, dtype=np.float32)
        weight = None if weight is None else np.ascontiguousarray(weight, dtype=np.float32)
        init_score = None if init_score is None else np.ascontiguousarray(init_score, dtype=np.float64)
        query = None if query is None else np.ascontiguousarray(query, dtype=np.int32)
        _check(lib().ORC_DatasetPushRowsWithMetadata(self.h, data.ctypes.data_as(C.c_void_p), C.c_int(1), C.c_int32(data.shape[0]),
                                                     C.c_int32(data.shape[1]), C.c_int32(start_row), _ptr(label, C.c_float),
                                                     _ptr(weight, C.c_float), _ptr(init_score, C.c_double), _ptr(query, C.c_int32),
                                                     C.c_int32(0)))

    def mark_finished(self):
        _check(lib().ORC_DatasetMarkFinished(self.h))

    def num_data(self):
        o = C.c_int()
        _check(lib().ORC_DatasetGetNumData(self.h, C.byref(o)))
        return o.value

    def num_feature(self):
        o = C.c_int()
        _check(lib().ORC_DatasetGetNumFeature(self.h, C.byref(o)))
        return o.value

    def num_used_feature(self):
        o = C.c_int()
        _check(lib().ORC_DatasetGetNumUsedFeature(self.h, C.byref(o)))
        return o.value

    def feature_info(self, real_feature):
        nb, mt, db, mf, tr = (C.c_int() for _ in range(5))
        mn, mx = C.c_double(), C.c_double()
        _check(lib().ORC_DatasetGetFeatureInfo(self.h, C.c_int(real_feature), C.byref(nb), C.byref(mt), C.byref(db), C.byref(mf),
                                               C.byref(tr), C.byref(mn), C.byref(mx)))
        return dict(num_bin=nb.value, missing_type=mt.value, default_bin=db.value, most_freq_bin=mf.value,
                    is_trivial=bool(tr.value), min_val=mn.value, max_val=mx.value)

    def bin_upper_bounds(self, real_feature):
        nb = self.feature_info(real_feature)["num_bin"]
        out = np.zeros(nb, dtype=np.float64)
        _check(lib().ORC_DatasetGetBinUpperBounds(self.h, C.c_int(real_feature), _ptr(out, C.c_double)))
        return out

    def bin_column(self, inner_feature):
        out = np.zeros(self.num_data(), dtype=np.uint16)
        _check(lib().ORC_DatasetGetBinColumn(self.h, C.c_int(inner_feature), _ptr(out, C.c_uint16)))
        return out

    def free(self):
        if self.h:
            lib().ORC_DatasetFree(self.h)
            self.h = None


class Booster:
    def __init__(self, datasets, params):
        if isinstance(datasets, Dataset):
            datasets = [datasets]
        self.datasets = datasets
        arr = (C.c_void_p * len(datasets))(*[d.h for d in datasets])
        out = C.c_void_p()
        _check(lib().ORC_BoosterCreate(arr, C.c_int(len(datasets)), params.encode(), C.byref(out)))
        self.h = out

    def update_one_iter(self):
        fin = C.c_int()
        _check(lib().ORC_BoosterUpdateOneIter(self.h, C.byref(fin)))
        return bool(fin.value)

    def update_one_iter_custom(self, grad, hess):
        g = np.ascontiguousarray(grad, dtype=np.float32)
        h = np.ascontiguousarray(hess, dtype=np.float32)
        fin = C.c_int()
        _check(lib().ORC_BoosterUpdateOneIterCustom(self.h, _ptr(g, C.c_float), _ptr(h, C.c_float), C.byref(fin)))
        return bool(fin.value)

    def save_model_to_string(self):
        n = C.c_int64()
        buf = C.create_string_buffer(1 << 16)
        _check(lib().ORC_BoosterSaveModelToString(self.h, C.c_int64(len(buf)), C.byref(n), buf))
        if n.value > len(buf):
            buf = C.create_string_buffer(n.value)
            _check(lib().ORC_BoosterSaveModelToString(self.h, C.c_int64(len(buf)), C.byref(n), buf))
        return buf.value.decode()

    def num_class(self):
        o = C.c_int()
        _check(lib().ORC_BoosterNumClass(self.h, C.byref(o)))
        return o.value

    def scores(self, shard=0):
        """Training scores of one shard; multiclass: class-major [num_class * n]."""
        out = np.zeros(self.datasets[shard].num_data() * self.num_class(), dtype=np.float64)
        _check(lib().ORC_BoosterGetScores(self.h, C.c_int(shard), _ptr(out, C.c_double)))
        return out

    def gradients(self, shard=0):
        n = self.datasets[shard].num_data() * self.num_class()
        g, h = np.zeros(n, dtype=np.float32), np.zeros(n, dtype=np.float32)
        _check(lib().ORC_BoosterGetGradients(self.h, C.c_int(shard), _ptr(g, C.c_float), _ptr(h, C.c_float)))
        return g, h

    def quant_gradients(self, shard=0):
        n = self.datasets[shard].num_data()
        g, h = np.zeros(n, dtype=np.int8), np.zeros(n, dtype=np.int8)
        sc = np.zeros(2, dtype=np.float64)
        _check(lib().ORC_BoosterGetQuantGradients(self.h, C.c_int(shard), _ptr(g, C.c_int8), _ptr(h, C.c_int8), _ptr(sc, C.c_double)))
        return g, h, sc

    def num_trees(self):
        o = C.c_int()
        _check(lib().ORC_BoosterNumTrees(self.h, C.byref(o)))
        return o.value

    def predict_raw(self, data):
        data = np.ascontiguousarray(data, dtype=np.float64)
        out = np.zeros(data.shape[0], dtype=np.float64)
        _check(lib().ORC_BoosterPredictRaw(self.h, _ptr(data, C.c_double), C.c_int(data.shape[0]), C.c_int(data.shape[1]),
                                           _ptr(out, C.c_double)))
        return out

    def free(self):
        if self.h:
            lib().ORC_BoosterFree(self.h)
            self.h = None


def set_num_threads(n):
    lib().ORC_SetNumThreads(C.c_int(n))


def max_threads():
    return lib().ORC_GetMaxThreads()


def fit(X, y, params, dataset_params=None, n_shards=1, num_iterations=None, weight=None, query=None, sample_rows=200000):
    """Convenience: the whole reference call sequence (§3.1-3.2 of SURVEY.md) on in-memory arrays, split into
    n_shards contiguous row shards.  Returns (booster, datasets)."""
    n = X.shape[0]
    sample = X[: min(sample_rows, n)]
    dparams = dataset_params if dataset_params is not None else params
    ref = Dataset.from_sample(sample, 1, n, dparams)
    bounds = [n * s // n_shards for s in range(n_shards + 1)]
    if query is not None and n_shards > 1:      # snap shard boundaries to query boundaries
        qb = np.flatnonzero(np.diff(query) != 0) + 1
        bounds = [0] + [int(qb[np.searchsorted(qb, b)]) if 0 < b < n else b for b in bounds[1:-1]] + [n]
    dss = []
    for s in range(n_shards):
        b, e = bounds[s], bounds[s + 1]
        d = ref.create_by_reference(e - b)
        d.init_streaming(weight is not None, False, query is not None)
        d.push_rows(X[b:e], 0, y[b:e], None if weight is None else weight[b:e], None, None if query is None else query[b:e])
        d.mark_finished()
        dss.append(d)
    bst = Booster(dss, params)
    if num_iterations is not None:
        for _ in range(num_iterations):
            if bst.update_one_iter():
                break
    return bst, dss
