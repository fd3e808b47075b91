"""Regression objectives beyond L2 (LightGBMRegressor's `objective`, `alpha`, `tweedieVariancePower`
params, RegressorTrainParams.scala:24-29): huber, fair, poisson, gamma, tweedie gradients on the GPU against the oracle's
restatement of upstream's regression_objective.hpp; the log-link objectives predict exp(raw score)."""
import ctypes as C

import numpy as np
import pytest

from helpers import NativeBooster, NativeDataset, strip_parameters
from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.estimators import LightGBMRegressor, LocalDataFrame
from synapseml_b200.lightgbm.model_text import TextModel
from synapseml_b200.lightgbm.native import lightgbmlib, validate
from test_gpu_predict import predict
from test_gpu_train import FLOAT_RTOL, train_both

pytestmark = pytest.mark.gpu

OBJECTIVES = {
    "huber": "objective=huber alpha=0.7",
    "fair": "objective=fair fair_c=2.0",
    "poisson": "objective=poisson",
    "gamma": "objective=gamma",
    "tweedie": "objective=tweedie tweedie_variance_power=1.3",
}


def count_data(n, positive):
    X, _ = S.dense100(n, f=12)
    rng = np.random.Generator(np.random.PCG64(11))
    mu = np.exp(0.4 * X[:, 0] - 0.3 * X[:, 1] + 0.2 * X[:, 2] * X[:, 3])
    y = rng.poisson(mu).astype(np.float32)
    if positive:
        y = (y + rng.uniform(0.1, 1.0, n)).astype(np.float32)
    return X, y


@pytest.mark.parametrize("name", sorted(OBJECTIVES))
def test_quantized_models_are_byte_identical(native, oracle, name):
    X, y = count_data(30000, positive=(name == "gamma"))
    params = OBJECTIVES[name] + " num_leaves=31 max_bin=63 learning_rate=0.1 use_quantized_grad=true stochastic_rounding=false verbosity=-1"
    nd, nb, ob = train_both(oracle, X, y, params, 10)
    ms = nb.model()
    assert strip_parameters(ms) == strip_parameters(ob.save_model_to_string())
    assert np.array_equal(nb.scores(), ob.scores())
    assert ("objective=%s\n" % name) in ms
    raw, normal = predict(nb.handle, X[:2000], 1), predict(nb.handle, X[:2000], 0)
    assert np.array_equal(raw, TextModel(ms).predict_raw(X[:2000]))
    if name in ("poisson", "gamma", "tweedie"):
        assert np.allclose(normal, np.exp(raw), rtol=1e-13, atol=0)
    else:
        assert np.array_equal(normal, raw)
    nb.free()
    nd.free()


def test_float_mode_poisson_with_weights(native, oracle):
    X, y = count_data(40000, positive=False)
    w = np.random.Generator(np.random.PCG64(4)).uniform(0.5, 2.0, len(y)).astype(np.float32)
    params = "objective=poisson num_leaves=31 max_bin=255 verbosity=-1"
    nd, nb, ob = train_both(oracle, X, y, params, 8, weight=w)
    sn, so = TextModel(nb.model()).structure(), TextModel(ob.save_model_to_string()).structure()
    assert sn[:3] == so[:3]
    a, b = nb.scores(), ob.scores()
    assert np.abs(a - b).max() <= FLOAT_RTOL * np.abs(b).max()
    nb.free()
    nd.free()


def test_default_metric_follows_the_objective(native):
    X, y = count_data(20000, positive=False)
    params = "objective=poisson num_leaves=15 max_bin=63 verbosity=-1"
    nd = NativeDataset(X, y, params)
    nb = NativeBooster(nd, params)
    for _ in range(5):
        nb.update()
    lib = lightgbmlib()
    n = C.c_int()
    out = np.zeros(4)
    validate(lib.LGBM_BoosterGetEval(nb.handle, 0, C.byref(n), out.ctypes.data_as(C.POINTER(C.c_double))))
    pred = np.maximum(np.exp(nb.scores()), 1e-10)
    assert n.value == 1 and np.isclose(out[0], np.mean(pred - y * np.log(pred)), rtol=1e-10)
    with pytest.raises(Exception):
        NativeBooster(NativeDataset(X, -y - 1, params), params)          # negative labels are refused
    nb.free()
    nd.free()


def test_gamma_deviance_metric(native):
    """metric=gamma_deviance (named by the reference's metric doc, LightGBMParams.scala:593): 2 * sum(w * (y / p - log(y / p) - 1))."""
    X, y = count_data(20000, positive=True)
    params = "objective=gamma metric=gamma,gamma_deviance num_leaves=15 max_bin=63 verbosity=-1"
    nd = NativeDataset(X, y, params)
    nb = NativeBooster(nd, params)
    for _ in range(5):
        nb.update()
    n = C.c_int()
    out = np.zeros(4)
    validate(lightgbmlib().LGBM_BoosterGetEval(nb.handle, 0, C.byref(n), out.ctypes.data_as(C.POINTER(C.c_double))))
    t = y / (np.exp(nb.scores()) + 1e-9)
    assert n.value == 2 and np.isclose(out[1], 2 * np.sum(t - np.log(t) - 1), rtol=1e-10)
    nb.free()
    nd.free()


def test_cross_entropy_objective_and_metric(native, oracle):
    """objective=cross_entropy (labels are probabilities; upstream xentropy_objective.hpp): byte-identical quantized model, sigmoid
    output, metric = -(y log p + (1 - y) log(1 - p)) averaged."""
    from helpers import oracle_dataset, strip_parameters
    X, _ = count_data(20000, positive=False)
    rng = np.random.Generator(np.random.PCG64(8))
    y = np.clip(1.0 / (1.0 + np.exp(-X[:, 0])) + rng.normal(0, 0.1, len(X)), 0, 1).astype(np.float32)
    params = "objective=cross_entropy num_leaves=15 max_bin=63 use_quantized_grad=true stochastic_rounding=false verbosity=-1"
    nd = NativeDataset(X, y, params)
    nb = NativeBooster(nd, params)
    _, dss, _ = oracle_dataset(oracle, X, y, params)
    ob = oracle.Booster(dss, params)
    for _ in range(6):
        assert nb.update() == ob.update_one_iter()
    ms = nb.model()
    assert strip_parameters(ms) == strip_parameters(ob.save_model_to_string()) and "objective=cross_entropy" in ms
    n = C.c_int()
    out = np.zeros(4)
    validate(lightgbmlib().LGBM_BoosterGetEval(nb.handle, 0, C.byref(n), out.ctypes.data_as(C.POINTER(C.c_double))))
    p = 1.0 / (1.0 + np.exp(-nb.scores()))
    assert n.value == 1 and np.isclose(out[0], -np.mean(y * np.log(p) + (1 - y) * np.log(1 - p)), rtol=1e-10)
    from synapseml_b200.lightgbm.booster import LightGBMBooster
    lb = LightGBMBooster(model_str=ms)
    assert np.allclose(lb.predict_for_mat(X[:500], LightGBMBooster.PREDICT_NORMAL), p[:500], rtol=1e-12)
    lb.free_native_memory()
    nb.free()
    nd.free()


def test_regressor_estimator_with_tweedie(native):
    X, y = count_data(30000, positive=False)
    df = LocalDataFrame.from_arrays(X, y, num_partitions=2)
    model = LightGBMRegressor(objective="tweedie", tweedieVariancePower=1.2, numIterations=30, numLeaves=15).fit(df)
    assert "objective=tweedie" in model.getNativeModel() and "[tweedie_variance_power: 1.2]" in model.getNativeModel()
    pred = model.transform(X)["prediction"]
    assert (pred > 0).all()                                           # exp link
    assert np.corrcoef(pred, y)[0, 1] > 0.5 and abs(pred.mean() - y.mean()) < 0.05 * y.mean()


# ---- l1 / quantile / mape: the objectives whose leaf outputs are re-fit as percentiles of the residuals (RenewTreeOutput)
RENEW = {
    "regression_l1": "objective=regression_l1",
    "quantile": "objective=quantile alpha=0.8",
    "mape": "objective=mape",
}


def skewed_targets(n):
    X, y = S.dense100(n, f=12)
    rng = np.random.Generator(np.random.PCG64(17))
    y = (y + np.abs(y) * 0.4 + rng.exponential(0.5, n)).astype(np.float32)      # continuous, asymmetric: percentiles differ from means
    return X, y


@pytest.mark.parametrize("weighted", [False, True])
@pytest.mark.parametrize("name", sorted(RENEW))
def test_percentile_objectives_quantized_models_are_byte_identical(native, oracle, name, weighted):
    X, y = skewed_targets(30000)
    w = np.random.Generator(np.random.PCG64(5)).uniform(0.5, 3.0, len(y)).astype(np.float32) if weighted else None
    params = RENEW[name] + " num_leaves=31 max_bin=63 learning_rate=0.2 use_quantized_grad=true stochastic_rounding=false verbosity=-1"
    nd, nb, ob = train_both(oracle, X, y, params, 8, weight=w)
    ms = nb.model()
    assert strip_parameters(ms) == strip_parameters(ob.save_model_to_string())
    assert np.array_equal(nb.scores(), ob.scores())
    assert ("objective=%s\n" % name) in ms
    raw = predict(nb.handle, X[:2000], 1)
    assert np.array_equal(raw, TextModel(ms).predict_raw(X[:2000])) and np.array_equal(predict(nb.handle, X[:2000], 0), raw)
    nb.free()
    nd.free()


def test_quantile_regressor_covers_alpha_of_the_labels_and_reports_its_metric(native):
    X, y = skewed_targets(60000)
    model_hi = LightGBMRegressor(objective="quantile", alpha=0.9, numIterations=60, numLeaves=31, learningRate=0.2).fit(LocalDataFrame.from_arrays(X, y))
    model_lo = LightGBMRegressor(objective="quantile", alpha=0.1, numIterations=60, numLeaves=31, learningRate=0.2).fit(LocalDataFrame.from_arrays(X, y))
    hi, lo = model_hi.transform(X)["prediction"], model_lo.transform(X)["prediction"]
    assert abs((y <= hi).mean() - 0.9) < 0.02 and abs((y <= lo).mean() - 0.1) < 0.02
    # the default metric of the objective, on the training data, against numpy
    params = "objective=quantile alpha=0.8 num_leaves=15 max_bin=63 verbosity=-1"
    nd = NativeDataset(X, y, params)
    nb = NativeBooster(nd, params)
    for _ in range(5):
        nb.update()
    n, out = C.c_int(), np.zeros(4)
    validate(lightgbmlib().LGBM_BoosterGetEval(nb.handle, 0, C.byref(n), out.ctypes.data_as(C.POINTER(C.c_double))))
    d = y.astype(np.float64) - nb.scores()
    assert n.value == 1 and np.isclose(out[0], np.where(d < 0, (np.float32(0.8) - 1.0) * d, 0.8 * d).mean(), rtol=1e-6)
    nb.free()
    nd.free()


def test_l1_float_mode_first_trees_and_scores(native, oracle):
    X, y = skewed_targets(40000)
    params = "objective=regression_l1 num_leaves=31 max_bin=255 learning_rate=0.2 verbosity=-1"
    nd, nb, ob = train_both(oracle, X, y, params, 8)
    sn, so = TextModel(nb.model()).structure(), TextModel(ob.save_model_to_string()).structure()
    assert sn[:3] == so[:3]
    a, b = nb.scores(), ob.scores()
    assert np.abs(a - b).max() <= FLOAT_RTOL * np.abs(b).max()
    nb.free()
    nd.free()
