"""Validation data, metrics and early stopping through the C ABI: LGBM_BoosterAddValidData / GetEvalNames /
GetEval against numpy / scikit-learn on the same scores, and the reference's early-stopping bookkeeping (TrainUtils.scala:137-169)."""
import ctypes as C

import numpy as np
import pytest

from helpers import NativeBooster, NativeDataset
from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LocalDataFrame
from synapseml_b200.lightgbm.native import lightgbmlib, validate

pytestmark = pytest.mark.gpu


def eval_names(h):
    lib = lightgbmlib()
    n = C.c_int()
    validate(lib.LGBM_BoosterGetEvalCounts(h, C.byref(n)))
    bufs = [C.create_string_buffer(64) for _ in range(n.value)]
    arr = (C.c_char_p * n.value)(*[C.cast(b, C.c_char_p) for b in bufs])
    got, need = C.c_int(), C.c_size_t()
    validate(lib.LGBM_BoosterGetEvalNames(h, n.value, C.byref(got), C.c_size_t(64), C.byref(need), arr))
    assert got.value == n.value and need.value <= 64
    return [b.value.decode() for b in bufs]


def get_eval(h, idx, k):
    out, n = np.zeros(k), C.c_int()
    validate(lightgbmlib().LGBM_BoosterGetEval(h, idx, C.byref(n), out.ctypes.data_as(C.POINTER(C.c_double))))
    assert n.value == k
    return out


def raw_scores(h, X):
    out, n = np.zeros(len(X)), C.c_int64()
    x = np.ascontiguousarray(X)
    validate(lightgbmlib().LGBM_BoosterPredictForMat(h, x.ctypes.data_as(C.c_void_p), 1, len(X), X.shape[1], 1, 1, 0, -1, b"", C.byref(n),
                                                     out.ctypes.data_as(C.POINTER(C.c_double))))
    return out


def test_binary_metrics_with_weights_on_training_and_validation_data(native):
    from sklearn.metrics import log_loss, roc_auc_score
    X, y = S.higgs_shape(60000)
    rng = np.random.Generator(np.random.PCG64(3))
    w = rng.uniform(0.5, 2.0, len(y)).astype(np.float32)
    Xv, yv, wv = X[40000:], y[40000:], w[40000:]
    X, y, w = X[:40000], y[:40000], w[:40000]
    Xv = Xv.copy(); Xv[:, 3] = np.round(Xv[:, 3])      # plenty of tied scores are produced by few-valued features anyway
    params = "objective=binary num_leaves=15 max_bin=63 metric=binary_logloss,auc,binary_error verbosity=-1"
    nd = NativeDataset(X, y, params, weight=w, micro_batches=[8192])
    vd = NativeDataset(Xv, yv, params, weight=wv, serialized=nd.serialized)
    nb = NativeBooster(nd, params)
    validate(nb.lib.LGBM_BoosterAddValidData(nb.handle, vd.handle))
    assert eval_names(nb.handle) == ["binary_logloss", "auc", "binary_error"]
    for it in range(6):
        nb.update()
        for idx, (xx, yy, ww) in enumerate([(X, y, w), (Xv, yv, wv)]):
            s = nb.scores() if idx == 0 else raw_scores(nb.handle, xx)
            p = 1.0 / (1.0 + np.exp(-s))
            got = get_eval(nb.handle, idx, 3)
            assert abs(got[0] - log_loss(yy, p, sample_weight=ww)) < 1e-12
            assert abs(got[1] - roc_auc_score(yy, s, sample_weight=ww)) < 1e-10
            assert abs(got[2] - np.average((p > 0.5) != (yy > 0), weights=ww)) < 1e-12
    # validation scores are advanced by the training loop: they equal a fresh prediction
    # (checked above through the metrics); out-of-range index fails loudly
    out, n = np.zeros(3), C.c_int()
    assert nb.lib.LGBM_BoosterGetEval(nb.handle, 2, C.byref(n), out.ctypes.data_as(C.POINTER(C.c_double))) == -1
    nb.free(); vd.free(); nd.free()


def test_regression_metrics(native):
    X, y = S.dense100(50000, f=20)
    params = "objective=regression num_leaves=31 max_bin=63 metric=l2,rmse verbosity=-1"
    nd = NativeDataset(X[:35000], y[:35000], params)
    vd = NativeDataset(X[35000:], y[35000:], params, serialized=nd.serialized)
    nb = NativeBooster(nd, params)
    validate(nb.lib.LGBM_BoosterAddValidData(nb.handle, vd.handle))
    assert eval_names(nb.handle) == ["l2", "rmse"]
    for _ in range(4):
        nb.update()
    s = raw_scores(nb.handle, X[35000:])
    mse = np.mean((s - y[35000:].astype(np.float64)) ** 2)
    got = get_eval(nb.handle, 1, 2)
    assert abs(got[0] - mse) < 1e-12 * max(1.0, mse) and abs(got[1] - np.sqrt(mse)) < 1e-12 * max(1.0, mse)
    got0 = get_eval(nb.handle, 0, 2)
    assert abs(got0[0] - np.mean((nb.scores() - y[:35000].astype(np.float64)) ** 2)) < 1e-12 * max(1.0, got0[0])
    nb.free(); vd.free(); nd.free()


def ndcg_ref(scores, labels, qb, ks):
    out = np.zeros(len(ks))
    for q in range(len(qb) - 1):
        s, l = scores[qb[q]:qb[q + 1]], labels[qb[q]:qb[q + 1]].astype(int)
        order = np.argsort(-s, kind="stable")
        gains = (2.0 ** l - 1.0)
        disc = 1.0 / np.log2(2.0 + np.arange(len(s)))
        ideal = np.sort(gains)[::-1]
        for j, k in enumerate(ks):
            kk = min(k, len(s))
            mx = np.sum(ideal[:kk] * disc[:kk])
            out[j] += 1.0 if mx <= 0 else np.sum(gains[order][:kk] * disc[:kk]) / mx
    return out / (len(qb) - 1)


def test_ndcg_at_k(native):
    X, y, q = S.mslr_shape(30000)
    params = "objective=lambdarank num_leaves=15 max_bin=63 eval_at=1,3,5 verbosity=-1"
    nd = NativeDataset(X, y, params, query=q)
    nb = NativeBooster(nd, params)
    assert eval_names(nb.handle) == ["ndcg@1", "ndcg@3", "ndcg@5"]
    qb = np.concatenate([[0], np.flatnonzero(np.diff(q)) + 1, [len(q)]])
    for _ in range(3):
        nb.update()
        got = get_eval(nb.handle, 0, 3)
        assert np.allclose(got, ndcg_ref(nb.scores(), y, qb, [1, 3, 5]), rtol=0, atol=1e-12)
    nb.free(); nd.free()


def map_ref(score, label, qb, ks):
    """MapMetric::CalMapAtK written out: sequential, per query, positions in stable order of falling score."""
    out = np.zeros(len(ks))
    for a, b in zip(qb[:-1], qb[1:]):
        s, rel = score[a:b], label[a:b] > 0.5
        order = np.argsort(-s, kind="stable")
        npos = int(rel.sum())
        hits, sum_ap, left = 0, 0.0, 0
        for j, k in enumerate(ks):
            kk = min(k, len(s))
            for pos in range(left, kk):
                if rel[order[pos]]:
                    hits += 1
                    sum_ap += hits / (pos + 1.0)
            out[j] += sum_ap / min(npos, kk) if npos > 0 else 1.0
            left = kk
    return out / (len(qb) - 1)


def test_map_at_k(native):
    """metric=map (LightGBMParams.scala:596 names it for ranking): mean average precision at eval_at, next to ndcg."""
    X, y, q = S.mslr_shape(30000)
    params = "objective=lambdarank num_leaves=15 max_bin=63 eval_at=1,3,10 metric=ndcg,map verbosity=-1"
    nd = NativeDataset(X, y, params, query=q)
    nb = NativeBooster(nd, params)
    assert eval_names(nb.handle) == ["ndcg@1", "ndcg@3", "ndcg@10", "map@1", "map@3", "map@10"]
    qb = np.concatenate([[0], np.flatnonzero(np.diff(q)) + 1, [len(q)]])
    for _ in range(3):
        nb.update()
        got = get_eval(nb.handle, 0, 6)
        assert np.allclose(got[:3], ndcg_ref(nb.scores(), y, qb, [1, 3, 10]), rtol=0, atol=1e-12)
        assert np.allclose(got[3:], map_ref(nb.scores(), y, qb, [1, 3, 10]), rtol=0, atol=1e-12)
    nb.free(); nd.free()


def test_early_stopping_through_the_estimator(native):
    # VerifyLightGBMClassifierStream.scala: validationIndicatorCol + earlyStoppingRound; noisy labels make validation loss turn early
    X, y = S.higgs_shape(30000)
    rng = np.random.Generator(np.random.PCG64(9))
    y = np.where(rng.random(len(y)) < 0.35, 1 - y, y).astype(np.float32)
    valid = np.zeros(len(y), dtype=bool)
    valid[20000:] = True
    df = LocalDataFrame.from_arrays(X, y, num_partitions=3, isVal=valid)
    est = LightGBMClassifier(numIterations=200, numLeaves=63, learningRate=0.3, minDataInLeaf=5, validationIndicatorCol="isVal",
                             earlyStoppingRound=5, metric="binary_logloss", samplingMode="fixed", isProvideTrainingMetric=True)
    model = est.fit(df)
    best = est.best_iteration
    hist = est.performance_measures["evalHistory"]
    assert best is not None and best < 150
    assert model.getNativeModel().count("Tree=") == best + 1
    vloss = [h[1]["binary_logloss"] for h in hist if h[1] is not None]
    assert int(np.argmin(vloss)) == best and len(vloss) == best + 1 + 5
    tloss = [h[0]["binary_logloss"] for h in hist if h[0] is not None]
    assert all(b < a for a, b in zip(tloss, tloss[1:]))      # training loss keeps falling


def test_validation_metrics_are_the_same_in_bulk_and_streaming_mode(native):
    X, y = S.higgs_shape(20000)
    valid = np.zeros(len(y), dtype=bool)
    valid[15000:] = True
    df = LocalDataFrame.from_arrays(X, y, num_partitions=2, isVal=valid)
    hist = {}
    for mode in ("streaming", "bulk"):
        est = LightGBMClassifier(numIterations=5, numLeaves=15, validationIndicatorCol="isVal", earlyStoppingRound=100, metric="auc",
                                 samplingMode="fixed", dataTransferMode=mode, useQuantizedGrad=True, stochasticRounding=False)
        est.fit(df)
        hist[mode] = [h[1]["auc"] for h in est.performance_measures["evalHistory"]]
    assert hist["streaming"] == hist["bulk"] and len(hist["bulk"]) == 5

