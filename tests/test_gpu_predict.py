"""Scoring through the C ABI: LGBM_BoosterPredictForMat / ...SingleRow on the GPU against an independent
numpy traversal of the model text (synapseml_b200/lightgbm/model_text.py is test-side here) and against the training scores;
LGBM_BoosterLoadModelFromString round trip."""
import ctypes as C

import numpy as np
import pytest

from helpers import NativeBooster, NativeDataset, edge_case_matrix
from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.booster import LightGBMBooster
from synapseml_b200.lightgbm.model_text import TextModel
from synapseml_b200.lightgbm.native import lightgbmlib, validate

pytestmark = pytest.mark.gpu


def predict(handle, X, predict_type, start=0, num=-1, dtype=np.float64, row_major=True):
    lib = lightgbmlib()
    n, f = X.shape
    x = np.ascontiguousarray(X if row_major else X.T, dtype=dtype)
    need = C.c_int64()
    validate(lib.LGBM_BoosterCalcNumPredict(handle, n, predict_type, start, num, C.byref(need)))
    out = np.full(max(need.value, 1), np.nan)
    got = C.c_int64()
    validate(lib.LGBM_BoosterPredictForMat(handle, x.ctypes.data_as(C.c_void_p), 1 if dtype == np.float64 else 0, n, f, 1 if row_major else 0,
                                           predict_type, start, num, b"", C.byref(got), out.ctypes.data_as(C.POINTER(C.c_double))))
    assert got.value == need.value
    return out[:got.value]


@pytest.fixture(scope="module")
def trained():
    X = edge_case_matrix(20000)          # NaNs, zeros, few-valued and constant columns
    rng = np.random.Generator(np.random.PCG64(5))
    y = (np.nan_to_num(X[:, 0]) + 0.5 * np.nan_to_num(X[:, 1]) + rng.normal(0, 0.5, len(X)) > 0).astype(np.float32)
    params = "objective=binary num_leaves=31 max_bin=63 min_data_in_leaf=5 learning_rate=0.2 verbosity=-1"
    nd = NativeDataset(X, y, params, micro_batches=[4096])
    nb = NativeBooster(nd, params)
    for _ in range(12):
        nb.update()
    yield X, nb
    nb.free()
    nd.free()


def test_raw_prediction_equals_numpy_traversal_and_training_scores(native, trained):
    X, nb = trained
    tm = TextModel(nb.model())
    raw = predict(nb.handle, X, 1)
    assert np.array_equal(raw, tm.predict_raw(X))              # same leaves, same summation order: the same doubles
    assert np.allclose(raw, nb.scores(), rtol=0, atol=1e-12)   # training scores were built by the apply-tree kernel from the bins


def test_normal_leaf_subrange_layouts_and_single_row(native, trained):
    X, nb = trained
    tm = TextModel(nb.model())
    raw = predict(nb.handle, X, 1)
    prob = predict(nb.handle, X, 0)
    assert np.allclose(prob, 1.0 / (1.0 + np.exp(-raw)), rtol=1e-13, atol=0)
    leaves = predict(nb.handle, X, 2).reshape(len(X), -1)
    want = np.stack([t.predict_leaf(X) for t in tm.trees], axis=1)
    assert np.array_equal(leaves.astype(np.int64), want)
    # trees [3, 8)
    part = predict(nb.handle, X, 1, start=3, num=5)
    assert np.array_equal(part, sum(t.predict(X) for t in tm.trees[3:8]))
    # float32 input (values are rounded to float32 first) and column-major input
    X32 = X.astype(np.float32)
    assert np.array_equal(predict(nb.handle, X32, 1, dtype=np.float32), tm.predict_raw(X32.astype(np.float64)))
    assert np.array_equal(predict(nb.handle, X, 1, row_major=False), raw)
    # one row at a time
    lib = lightgbmlib()
    for i in (0, 17, 4999):
        out, n = np.zeros(1), C.c_int64()
        row = np.ascontiguousarray(X[i])
        validate(lib.LGBM_BoosterPredictForMatSingleRow(nb.handle, row.ctypes.data_as(C.c_void_p), 1, X.shape[1], 1, 1, 0, -1, b"", C.byref(n),
                                                        out.ctypes.data_as(C.POINTER(C.c_double))))
        assert n.value == 1 and out[0] == raw[i]
    # wrong width and an unknown predict type fail loudly
    bad = np.zeros((2, X.shape[1] + 1))
    n = C.c_int64()
    o = np.zeros(2)
    assert lib.LGBM_BoosterPredictForMat(nb.handle, bad.ctypes.data_as(C.c_void_p), 1, 2, bad.shape[1], 1, 1, 0, -1, b"", C.byref(n),
                                         o.ctypes.data_as(C.POINTER(C.c_double))) == -1
    assert lib.LGBM_BoosterPredictForMat(nb.handle, X.ctypes.data_as(C.c_void_p), 1, 2, X.shape[1], 1, 7, 0, -1, b"", C.byref(n),
                                         o.ctypes.data_as(C.POINTER(C.c_double))) == -1


def test_sparse_rows_score_like_dense_rows(native, trained):
    import scipy.sparse as sp
    X, nb = trained
    Xz = np.nan_to_num(X[:3000])
    M = sp.csr_matrix(Xz)
    indptr, indices, vals = M.indptr.astype(np.int32), M.indices.astype(np.int32), M.data.astype(np.float64)
    out, n = np.zeros(len(Xz)), C.c_int64()
    validate(lightgbmlib().LGBM_BoosterPredictForCSR(nb.handle, indptr.ctypes.data_as(C.c_void_p), 2, indices.ctypes.data_as(C.POINTER(C.c_int32)),
                                                     vals.ctypes.data_as(C.c_void_p), 1, C.c_int64(len(indptr)), C.c_int64(len(vals)),
                                                     C.c_int64(X.shape[1]), 1, 0, -1, b"", C.byref(n), out.ctypes.data_as(C.POINTER(C.c_double))))
    assert n.value == len(Xz) and np.array_equal(out, predict(nb.handle, Xz, 1))


def test_booster_rebuilt_from_the_model_string_predicts_identically(native, trained):
    X, nb = trained
    text = nb.model()
    lb = LightGBMBooster(model_str=text)
    assert lb.num_iterations == 12 and lb.num_total_model() == 12 and lb.num_features() == X.shape[1]
    assert lb.save_to_string() == text
    assert np.array_equal(lb.predict_for_mat(X, LightGBMBooster.PREDICT_RAW), predict(nb.handle, X, 1))
    assert lb.score(X[3], raw=True) == predict(nb.handle, X[3:4], 1)[0]
    lb.free_native_memory()


def test_regression_and_large_batches(native):
    X, y = S.dense100(300000, f=20)
    params = "objective=regression num_leaves=63 max_bin=255 verbosity=-1"
    nd = NativeDataset(X, y, params, micro_batches=[1 << 16])
    nb = NativeBooster(nd, params)
    for _ in range(5):
        nb.update()
    tm = TextModel(nb.model())
    raw = predict(nb.handle, X, 1)
    assert np.array_equal(raw, tm.predict_raw(X))
    assert np.array_equal(predict(nb.handle, X, 0), raw)       # regression: no output transform
    nb.free()
    nd.free()


def test_concurrent_scoring_on_one_handle(native):
    """The reference scores from many task threads on ONE booster handle (LightGBMBooster.scala:67-140: thread-local output buffers,
    shared boosterPtr): concurrent calls must each get the predictions of their own rows."""
    import threading
    from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LocalDataFrame
    X, y = S.higgs_shape(30000)
    model = LightGBMClassifier(numIterations=10, numLeaves=15, samplingMode="fixed", microBatchSize=4096).fit(LocalDataFrame.from_arrays(X, y))
    parts = [X[i * 3000:(i + 1) * 3000 + 17 * i] for i in range(8)]
    want = [model.predict_raw(p) for p in parts]
    got = [None] * len(parts)
    errors = []

    def work(i):
        try:
            for _ in range(5):
                got[i] = model.predict_raw(parts[i])
                assert np.array_equal(got[i], want[i])
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=work, args=(i,)) for i in range(len(parts))]
    [t.start() for t in threads]
    [t.join() for t in threads]
    assert not errors, errors[0]
