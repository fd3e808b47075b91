"""Shared test helpers: replay the reference's native call sequence against the CUDA library through
its C ABI, and against the oracle."""
import ctypes as C

import numpy as np

from synapseml_b200.lightgbm.dataset import ReferenceDatasetUtils
from synapseml_b200.lightgbm.native import lightgbmlib, validate


def edge_case_matrix(n=6000, seed=3):
    """Columns exercising the bin finder: gaussian, exact zeros + positives, negatives + zeros + positives, NaNs, few distinct
    values, constant (trivial), tiny magnitudes around the zero threshold, integers with ties, heavy-tailed."""
    rng = np.random.Generator(np.random.PCG64(seed))
    cols = [
        rng.standard_normal(n),
        np.where(rng.random(n) < 0.6, 0.0, rng.exponential(1.0, n)),
        np.where(rng.random(n) < 0.3, 0.0, rng.standard_normal(n) * 3),
        np.where(rng.random(n) < 0.1, np.nan, rng.standard_normal(n)),
        rng.choice(np.array([0.0, 1.0865, 2.1731]), n),
        np.full(n, 4.25),
        rng.standard_normal(n) * 1e-36,
        rng.poisson(3.0, n).astype(np.float64),
        np.exp(rng.standard_normal(n) * 3),
        np.where(rng.random(n) < 0.5, np.nan, np.where(rng.random(n) < 0.5, 0.0, -rng.exponential(2.0, n))),
        rng.integers(-3, 4, n).astype(np.float64),
        np.where(rng.random(n) < 0.97, 0.0, 1.0),
    ]
    return np.stack(cols, axis=1)


class NativeDataset:
    """LGBM_* call sequence of the streaming path on in-memory arrays."""

    def __init__(self, X, y, params, num_dist_rows=None, sample_rows=200000, micro_batches=None, weight=None, query=None,
                 init_score=None, serialized=None, n_threads=1):
        lib = lightgbmlib()
        self.lib = lib
        n, f = X.shape
        self.n, self.f = n, f
        if serialized is None:
            serialized = ReferenceDatasetUtils.create_reference_dataset_from_sample(params, "features", num_dist_rows or n, f, X[:min(sample_rows, n)])
        self.serialized = serialized
        self.handle = ReferenceDatasetUtils.get_initialized_reference_dataset(serialized, n, 1, params, weight is not None,
                                                                              init_score is not None, query is not None, n_threads)
        if micro_batches is None:
            micro_batches = [n]
        y = np.ascontiguousarray(y, dtype=np.float32)
        s = 0
        k = 0
        while s < n:
            mb = micro_batches[k % len(micro_batches)]
            k += 1
            e = min(n, s + mb)
            self.push(X[s:e], s, y[s:e], None if weight is None else weight[s:e], None if init_score is None else init_score[s:e],
                      None if query is None else query[s:e], 0)
            s = e
        validate(lib.LGBM_DatasetMarkFinished(self.handle))

    def push(self, xb, start, yb, wb, ib, qb, tid):
        xb = np.ascontiguousarray(xb, dtype=np.float64)
        yb = np.ascontiguousarray(yb, dtype=np.float32)
        wb = None if wb is None else np.ascontiguousarray(wb, dtype=np.float32)
        ib = None if ib is None else np.ascontiguousarray(ib, dtype=np.float64)
        qb = None if qb is None else np.ascontiguousarray(qb, dtype=np.int32)
        validate(self.lib.LGBM_DatasetPushRowsWithMetadata(
            self.handle, xb.ctypes.data_as(C.c_void_p), 1, C.c_int32(xb.shape[0]), C.c_int32(self.f), C.c_int32(start),
            yb.ctypes.data_as(C.POINTER(C.c_float)), None if wb is None else wb.ctypes.data_as(C.POINTER(C.c_float)),
            None if ib is None else ib.ctypes.data_as(C.POINTER(C.c_double)), None if qb is None else qb.ctypes.data_as(C.POINTER(C.c_int32)),
            C.c_int32(tid)))

    def feature_info(self, j):
        nb, mt, db, mf, tr = (C.c_int() for _ in range(5))
        mn, mx = C.c_double(), C.c_double()
        validate(self.lib.LGBM_B200_DatasetGetFeatureInfo(self.handle, j, C.byref(nb), C.byref(mt), C.byref(db), C.byref(mf), C.byref(tr),
                                                          C.byref(mn), C.byref(mx)))
        return dict(num_bin=nb.value, missing_type=mt.value, default_bin=db.value, most_freq_bin=mf.value, is_trivial=bool(tr.value),
                    min_val=mn.value, max_val=mx.value)

    def bounds(self, j):
        ub = np.zeros(self.feature_info(j)["num_bin"])
        validate(self.lib.LGBM_B200_DatasetGetBinUpperBounds(self.handle, j, ub.ctypes.data_as(C.POINTER(C.c_double))))
        return ub

    def num_used(self):
        nu = C.c_int()
        validate(self.lib.LGBM_B200_DatasetGetNumUsedFeature(self.handle, C.byref(nu)))
        return nu.value

    def bin_column(self, j):
        col = np.zeros(self.n, dtype=np.uint16)
        validate(self.lib.LGBM_B200_DatasetCopyBinColumn16(self.handle, j, col.ctypes.data_as(C.POINTER(C.c_uint16))))
        return col if col.max(initial=0) > 255 else col.astype(np.uint8)

    def row_major(self, row=0, nrows=None):
        nrows = self.n - row if nrows is None else nrows
        stride = C.c_int()
        validate(self.lib.LGBM_B200_DatasetCopyRowMajor(self.handle, 0, 0, C.byref(stride), None))
        out = np.zeros((nrows, stride.value), dtype=np.uint8)
        validate(self.lib.LGBM_B200_DatasetCopyRowMajor(self.handle, C.c_int64(row), C.c_int64(nrows), C.byref(stride), out.ctypes.data_as(C.POINTER(C.c_uint8))))
        return out

    def free(self):
        if self.handle is not None:
            validate(self.lib.LGBM_DatasetFree(self.handle))
            self.handle = None


class NativeBooster:
    def __init__(self, ds, params):
        self.lib = lightgbmlib()
        self.ds = ds
        self.handle = C.c_void_p()
        validate(self.lib.LGBM_BoosterCreate(ds.handle, params.encode(), C.byref(self.handle)))

    def update(self):
        fin = C.c_int()
        validate(self.lib.LGBM_BoosterUpdateOneIter(self.handle, C.byref(fin)))
        return fin.value == 1

    def model(self):
        n = C.c_int64()
        sb = C.create_string_buffer(1 << 22)
        validate(self.lib.LGBM_BoosterSaveModelToString(self.handle, 0, -1, 0, C.c_int64(len(sb)), C.byref(n), sb))
        assert n.value <= len(sb)
        return sb.value.decode()

    def num_classes(self):
        k = C.c_int()
        validate(self.lib.LGBM_BoosterGetNumClasses(self.handle, C.byref(k)))
        return k.value

    def scores(self):
        sc = np.zeros(self.ds.n * self.num_classes())      # multiclass: class-major
        n64 = C.c_int64()
        validate(self.lib.LGBM_BoosterGetPredict(self.handle, 0, C.byref(n64), sc.ctypes.data_as(C.POINTER(C.c_double))))
        return sc

    def gradients(self):
        n = self.ds.n * self.num_classes()
        g, h = np.zeros(n, dtype=np.float32), np.zeros(n, dtype=np.float32)
        validate(self.lib.LGBM_B200_BoosterGetGradients(self.handle, g.ctypes.data_as(C.POINTER(C.c_float)), h.ctypes.data_as(C.POINTER(C.c_float))))
        return g, h

    def quant_gradients(self):
        g, h = np.zeros(self.ds.n, dtype=np.int8), np.zeros(self.ds.n, dtype=np.int8)
        sc = np.zeros(2)
        validate(self.lib.LGBM_B200_BoosterGetQuantGradients(self.handle, g.ctypes.data_as(C.POINTER(C.c_int8)), h.ctypes.data_as(C.POINTER(C.c_int8)),
                                                             sc.ctypes.data_as(C.POINTER(C.c_double))))
        return g, h, sc

    def root_histogram(self):
        out = np.zeros((self.ds.num_used(), 256, 2), dtype=np.int64)
        validate(self.lib.LGBM_B200_BoosterRootHistogram(self.handle, out.ctypes.data_as(C.POINTER(C.c_int64))))
        return out

    def free(self):
        if self.handle is not None:
            validate(self.lib.LGBM_BoosterFree(self.handle))
            self.handle = None


def strip_parameters(model_str):
    """The parameters block echoes the received param string; the parity artefact is everything before it."""
    return model_str.split("parameters:")[0]


def oracle_dataset(O, X, y, params, num_dist_rows=None, sample_rows=200000, weight=None, query=None, n_shards=1):
    n = X.shape[0]
    ref = O.Dataset.from_sample(X[:min(sample_rows, n)], 1, num_dist_rows or n, params)
    bounds = [n * s // n_shards for s in range(n_shards + 1)]
    if query is not None and n_shards > 1:
        qb = np.flatnonzero(np.diff(query) != 0) + 1
        bounds = [0] + [int(qb[min(np.searchsorted(qb, b), len(qb) - 1)]) for b in bounds[1:-1]] + [n]
    dss = []
    for s in range(n_shards):
        b, e = bounds[s], bounds[s + 1]
        d = ref.create_by_reference(e - b)
        d.init_streaming(weight is not None, False, query is not None)
        d.push_rows(X[b:e], 0, y[b:e], None if weight is None else weight[b:e], None, None if query is None else query[b:e])
        d.mark_finished()
        dss.append(d)
    return ref, dss, bounds
