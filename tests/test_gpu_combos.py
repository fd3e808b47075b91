"""Seeded random combinations of the learner's options — objective, boosting type, categorical columns, bin width (8-bit and wide
features), missing-value handling, monotone constraints, column / row sampling, regularisation, weights — each trained for a few
iterations in quantized mode on the GPU and by the oracle: the models must be byte-identical.  The single-option tests pin each
feature; this one pins their interactions (categorical + zero_as_missing, wide features + constraints + sampling, rf / dart / goss on
categorical data, ...)."""
import ctypes as C

import numpy as np
import pytest

from helpers import NativeBooster, NativeDataset, oracle_dataset, strip_parameters
from synapseml_b200.lightgbm.booster import LightGBMBooster
from synapseml_b200.lightgbm.model_text import TextModel
from synapseml_b200.lightgbm.native import lightgbmlib, validate

pytestmark = pytest.mark.gpu


def combo_data(n, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    X = rng.normal(size=(n, 10))
    X[:, 1] = rng.integers(0, 4, n)                                   # few categories
    X[:, 2] = np.where(rng.random(n) < 0.5, 0.0, rng.exponential(1.0, n))      # exact zeros
    X[:, 3] = rng.integers(0, 30, n)                                  # many categories
    X[rng.random(n) < 0.08, 4] = np.nan
    X[rng.random(n) < 0.04, 3] = np.nan
    X[:, 5] = rng.integers(-2, 3, n).astype(np.float64)               # few distinct values with zero
    X[:, 6] = 3.0                                                     # trivial
    eff = rng.normal(size=30)
    z = (X[:, 0] - 0.7 * X[:, 7] + 0.5 * np.sin(2 * X[:, 8]) + eff[np.nan_to_num(X[:, 3]).astype(int)] + (X[:, 1] == 2) * 0.8
         + 0.4 * X[:, 2] - 0.6 * np.nan_to_num(X[:, 4]) + 0.2 * X[:, 5])
    return X, z, rng


def random_combo(seed):
    r = np.random.Generator(np.random.PCG64(1000 + seed))
    pick = lambda *xs: xs[int(r.integers(0, len(xs)))]
    objective = pick("binary", "regression", "multiclass", "regression_l1", "multiclassova", "huber", "lambdarank", "poisson", "quantile", "fair",
                     "tweedie", "mape", "gamma", "rank_xendcg", "cross_entropy")
    boosting = pick("gbdt", "gbdt", "gbdt", "rf", "dart", "goss")
    p = ["objective=" + objective, "num_leaves=%d" % pick(7, 15, 31), "max_bin=%d" % pick(15, 63, 255, 600), "min_data_in_leaf=%d" % pick(5, 20, 60),
         "learning_rate=%s" % pick("0.1", "0.3")]
    if objective in ("multiclass", "multiclassova"):
        p.append("num_class=3")
    if pick(True, True, False):
        p.append("categorical_feature=" + pick("1,3", "3", "1"))
        p += pick([], ["max_cat_to_onehot=8"], ["cat_smooth=2", "min_data_per_group=20"], ["max_cat_threshold=4", "cat_l2=1"])
    if pick(True, False, False):
        p.append("zero_as_missing=true")
    elif pick(True, False, False, False):
        p.append("use_missing=false")
    if pick(True, False, False):
        p.append("monotone_constraints=1,0,0,0,0,0,0,-1,0,0")
        p += pick([], ["monotone_penalty=0.8"])
    bag = boosting == "rf" or pick(True, False, False)
    if boosting != "goss" and bag:
        p += ["bagging_freq=1", "bagging_fraction=%s" % pick("0.6", "0.8")]
    if boosting == "rf" or pick(True, False, False):
        p.append("feature_fraction=%s" % pick("0.6", "0.8"))
    if boosting not in ("rf",) and pick(True, False, False, False):
        p.append("feature_fraction_bynode=%s" % pick("0.5", "0.7"))
    if boosting == "dart":
        p += ["drop_rate=0.3", "skip_drop=0.2"]
    if boosting == "goss":
        p += ["top_rate=0.3", "other_rate=0.2"]
    p += pick([], ["lambda_l2=1.5"], ["lambda_l1=0.3", "lambda_l2=0.5"], ["min_gain_to_split=0.5"], ["max_delta_step=0.4"], ["max_depth=4"])
    p.append("boosting=" + boosting)
    p.append("stochastic_rounding=%s" % pick("false", "true"))
    p.append("seed=%d" % int(r.integers(1, 100)))
    weighted = bool(pick(True, False, False))
    return objective, " ".join(p) + " use_quantized_grad=true verbosity=-1", weighted


def labels_for(objective, z, rng):
    if objective in ("multiclass", "multiclassova"):
        return np.digitize(z + rng.normal(0, 0.3, len(z)), [-0.8, 0.9]).astype(np.float32)
    if objective == "binary":
        return (z + rng.normal(0, 0.5, len(z)) > 0).astype(np.float32)
    if objective in ("lambdarank", "rank_xendcg"):
        return np.digitize(z + rng.normal(0, 0.5, len(z)), [-0.5, 0.8, 2.0]).astype(np.float32)       # relevance 0..3
    if objective == "cross_entropy":
        return np.clip(1.0 / (1.0 + np.exp(-z)) + rng.normal(0, 0.1, len(z)), 0.0, 1.0).astype(np.float32)      # probabilities as labels
    if objective in ("poisson", "tweedie"):
        return rng.poisson(np.exp(np.clip(0.4 * z, -3, 2))).astype(np.float32)
    if objective == "gamma":
        return (rng.gamma(2.0, np.exp(np.clip(0.4 * z, -3, 2)) / 2.0) + 1e-3).astype(np.float32)
    return (z + rng.normal(0, 0.3, len(z))).astype(np.float32)


@pytest.mark.parametrize("seed", range(160))
def test_random_option_combinations_are_byte_identical(native, oracle, seed):
    objective, params, weighted = random_combo(seed)
    X, z, rng = combo_data(12000, seed)
    y = labels_for(objective, z, rng)
    w = rng.uniform(0.5, 2.0, len(y)).astype(np.float32) if weighted else None
    query = None
    if objective in ("lambdarank", "rank_xendcg"):       # queries of 1..60 documents, contiguous
        sizes = []
        while sum(sizes) < len(y):
            sizes.append(int(rng.integers(1, 61)))
        query = np.repeat(np.arange(len(sizes)), sizes)[:len(y)].astype(np.int32)
    nd = NativeDataset(X, y, params, weight=w, query=query, micro_batches=[4096, 100])
    nb = NativeBooster(nd, params)
    # a validation set (other rows of the same distribution, built on the training data's bin mappers): its scores are advanced by
    # the training loop through the tree walk over bins and must equal a fresh prediction on the raw values
    Xv, zv, rngv = combo_data(3000, 10_000 + seed)
    Xv[:200, 3] = 55
    if "zero_as_missing" not in params:
        # (under zero_as_missing a value below everything the sample held lands in the zero bin, i.e. counts as missing over bins but
        # not over values — upstream's own inconsistency between training-time validation scores and prediction; not probed here)
        Xv[200:300, 1] = -1
    vq = None if query is None else np.repeat(np.arange(100), 30).astype(np.int32)
    vd = NativeDataset(Xv, labels_for(objective, zv, rngv), params, query=vq, weight=None if w is None else np.ones(len(Xv), dtype=np.float32),
                       serialized=nd.serialized)
    validate(nb.lib.LGBM_BoosterAddValidData(nb.handle, vd.handle))
    _, dss, _ = oracle_dataset(oracle, X, y, params, weight=w, query=query)
    ob = oracle.Booster(dss, params)
    for _ in range(6):
        assert nb.update() == ob.update_one_iter(), params
    assert strip_parameters(nb.model()) == strip_parameters(ob.save_model_to_string()), params
    a, b = nb.scores(), ob.scores()
    if "boosting=dart" in params:
        assert np.allclose(a, b, rtol=0, atol=1e-12), params
    else:
        assert np.array_equal(a, b), params
    # scoring on raw values — unseen / negative / fractional categories, NaNs and zeros included — against an independent traversal
    # of the model text; leaf indices; feature contributions sum to the raw score; the model string survives a reload
    ms = nb.model()
    tm = TextModel(ms)
    Xp = X[:1500].copy()
    Xp[:100, 3] = 77; Xp[100:200, 3] = -4; Xp[200:300, 1] = 2.7; Xp[300:400, 0] = np.nan; Xp[400:500, 2] = 0.0; Xp[500:600, 4] = 0.0
    K = tm.num_tree_per_iteration
    raw = predict(nb.handle, Xp, 1).reshape(len(Xp), K)
    assert np.array_equal(raw, tm.predict_raw(Xp).reshape(len(Xp), K)), params
    leaves = predict(nb.handle, Xp, 2).reshape(len(Xp), -1)
    assert np.array_equal(leaves.astype(np.int64), np.stack([t.predict_leaf(Xp) for t in tm.trees], axis=1)), params
    contrib = predict(nb.handle, Xp[:200], 3).reshape(200, K, X.shape[1] + 1)
    assert np.allclose(contrib.sum(axis=2), raw[:200], rtol=0, atol=1e-9), params
    vs = np.zeros(len(Xv) * K)
    vn = C.c_int64()
    validate(nb.lib.LGBM_BoosterGetPredict(nb.handle, 1, C.byref(vn), vs.ctypes.data_as(C.POINTER(C.c_double))))
    assert vn.value == len(Xv) * K
    fresh = predict(nb.handle, Xv, 1).reshape(len(Xv), K)
    assert np.allclose(vs.reshape(K, len(Xv)).T, fresh, rtol=0, atol=1e-10), params      # (RF / DART rescale running sums: not bit-equal)
    lb = LightGBMBooster(model_str=ms)
    assert lb.save_to_string() == ms
    assert np.array_equal(lb.predict_for_mat(Xp, LightGBMBooster.PREDICT_RAW).reshape(len(Xp), K), raw), params
    lb.free_native_memory()
    nb.free()
    vd.free()
    nd.free()


def predict(handle, X, predict_type):
    lib = lightgbmlib()
    n, f = X.shape
    x = np.ascontiguousarray(X, dtype=np.float64)
    need = C.c_int64()
    validate(lib.LGBM_BoosterCalcNumPredict(handle, n, predict_type, 0, -1, C.byref(need)))
    out = np.full(max(need.value, 1), np.nan)
    got = C.c_int64()
    validate(lib.LGBM_BoosterPredictForMat(handle, x.ctypes.data_as(C.c_void_p), 1, n, f, 1, predict_type, 0, -1, b"", C.byref(got),
                                           out.ctypes.data_as(C.POINTER(C.c_double))))
    return out[:got.value]
