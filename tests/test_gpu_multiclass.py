"""Multiclass (objective=multiclass, K trees per iteration; LightGBMClassifier.scala:59-67): softmax gradients,
per-class trees, class-major scores, model text with num_class / num_tree_per_iteration, softmax scoring and the multi metrics,
against the oracle (quantized: byte-identical) and against numpy."""
import ctypes as C

import numpy as np
import pytest

from helpers import NativeBooster, NativeDataset, oracle_dataset, strip_parameters
from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.model_text import TextModel

pytestmark = pytest.mark.gpu


def three_class(n, seed=0):
    X, _ = S.higgs_shape(n)
    rng = np.random.Generator(np.random.PCG64(seed))
    z = X[:, 0] + 0.5 * X[:, 2] + 0.3 * rng.standard_normal(n)
    y = np.digitize(z, [-0.6, 0.7]).astype(np.float32)
    return X, y


def train_both(oracle, X, y, params, iters, weight=None):
    nd = NativeDataset(X, y, params, weight=weight, micro_batches=[8192])
    nb = NativeBooster(nd, params)
    _, dss, _ = oracle_dataset(oracle, X, y, params, weight=weight)
    ob = oracle.Booster(dss, params)
    for _ in range(iters):
        assert nb.update() == ob.update_one_iter()
    return nd, nb, ob


def test_first_iteration_gradients_are_bit_exact(native, oracle):
    X, y = three_class(30000)
    params = "objective=multiclass num_class=3 num_leaves=15 max_bin=63 min_data_in_leaf=20 verbosity=-1"
    nd, nb, ob = train_both(oracle, X, y, params, 1)
    g, h = nb.gradients()
    og, oh = ob.gradients()
    assert np.array_equal(g, og) and np.array_equal(h, oh)
    nb.free()
    nd.free()


@pytest.mark.parametrize("weighted", [False, True])
def test_quantized_multiclass_model_is_byte_identical(native, oracle, weighted):
    X, y = three_class(60000)
    w = np.random.Generator(np.random.PCG64(4)).uniform(0.5, 2.0, len(y)).astype(np.float32) if weighted else None
    params = ("objective=multiclass num_class=3 num_leaves=31 learning_rate=0.2 max_bin=63 min_data_in_leaf=20 use_quantized_grad=true "
              "stochastic_rounding=%s seed=3 verbosity=-1" % ("true" if weighted else "false"))
    nd, nb, ob = train_both(oracle, X, y, params, 8, weight=w)
    ms = nb.model()
    assert strip_parameters(ms) == strip_parameters(ob.save_model_to_string())
    assert "num_class=3" in ms and "num_tree_per_iteration=3" in ms and "objective=multiclass num_class:3" in ms
    sc = nb.scores()
    assert np.array_equal(sc, ob.scores())
    m = TextModel(ms)
    assert len(m.trees) == 24
    assert np.allclose(m.predict_raw(X[:5000]), sc.reshape(3, -1).T[:5000], rtol=0, atol=1e-12)      # class-major scores == tree traversal
    nb.free()
    nd.free()


def test_float_multiclass_first_trees_and_scores(native, oracle):
    X, y = three_class(60000)
    params = "objective=multiclass num_class=3 num_leaves=31 max_bin=255 min_data_in_leaf=20 verbosity=-1"
    nd, nb, ob = train_both(oracle, X, y, params, 6)
    a, b = TextModel(nb.model()), TextModel(ob.save_model_to_string())
    assert a.structure()[:6] == b.structure()[:6]
    sa, sb = nb.scores(), ob.scores()
    assert np.abs(sa - sb).max() <= 1e-5 * np.abs(sb).max()
    nb.free()
    nd.free()


def test_estimator_infers_num_class_and_scores_with_softmax(native):
    from synapseml_b200.lightgbm.booster import LightGBMBooster
    from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LocalDataFrame
    X, y = three_class(40000)
    est = LightGBMClassifier(objective="multiclass", numIterations=10, numLeaves=15, samplingMode="fixed", microBatchSize=4096,
                             isProvideTrainingMetric=True, metric="multi_logloss,multi_error")
    model = est.fit(LocalDataFrame.from_arrays(X, y, num_partitions=3))
    assert "num_class=3" in model.model_str
    out = model.transform(X[:8000])
    m = TextModel(model.model_str)
    assert out["rawPrediction"].shape == (8000, 3) and out["probability"].shape == (8000, 3)
    assert np.array_equal(out["rawPrediction"], m.predict_raw(X[:8000]))
    assert np.allclose(out["probability"], m.predict(X[:8000]), rtol=1e-12, atol=1e-15)
    assert np.allclose(out["probability"].sum(axis=1), 1.0)
    assert (out["prediction"] == y[:8000]).mean() > 0.85
    # the training metrics of the last iteration against numpy on the whole training set
    hist = est.performance_measures["evalHistory"]
    last = hist[-1][0]
    p = m.predict(X)
    yi = y.astype(int)
    logloss = -np.log(np.maximum(p[np.arange(len(y)), yi], 1e-15)).mean()
    raw = m.predict_raw(X)
    err = ((raw >= raw[np.arange(len(y)), yi][:, None]).sum(axis=1) > 1).mean()
    assert abs(last["multi_logloss"] - logloss) < 1e-10 and abs(last["multi_error"] - err) < 1e-12
    # leaf indices: one per tree
    leaves = model.predict_leaf(X[:100])
    assert leaves.size == 100 * len(m.trees)


@pytest.mark.parametrize("variant", ["plain", "weighted_unbalanced"])
def test_multiclassova_matches_the_oracle(native, oracle, variant):
    """objective=multiclassova (named by the reference's objective doc, LightGBMParams.scala:383-386): one binary log-loss problem per
    class (upstream MulticlassOVA), K trees per iteration."""
    X, y = three_class(60000)
    w = np.random.Generator(np.random.PCG64(4)).uniform(0.5, 2.0, len(y)).astype(np.float32) if variant != "plain" else None
    extra = " is_unbalance=true sigmoid=1.5" if variant != "plain" else ""
    params = ("objective=multiclassova num_class=3 num_leaves=31 learning_rate=0.2 max_bin=63 min_data_in_leaf=20 use_quantized_grad=true "
              "stochastic_rounding=false verbosity=-1" + extra)
    nd, nb, ob = train_both(oracle, X, y, params, 1, weight=w)
    g, h = nb.gradients()
    og, oh = ob.gradients()
    assert np.array_equal(g, og) and np.array_equal(h, oh)          # the first iteration's gradients of all three classes
    for _ in range(7):
        assert nb.update() == ob.update_one_iter()
    ms = nb.model()
    assert strip_parameters(ms) == strip_parameters(ob.save_model_to_string())
    assert "objective=multiclassova num_class:3 sigmoid:%s" % ("1.5" if extra else "1") in ms and "num_tree_per_iteration=3" in ms
    assert np.array_equal(nb.scores(), ob.scores())
    nb.free()
    nd.free()


def test_multiclassova_estimator_scores_with_one_sigmoid_per_class(native):
    from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LocalDataFrame
    X, y = three_class(40000)
    est = LightGBMClassifier(objective="multiclassova", numIterations=10, numLeaves=15, samplingMode="fixed", microBatchSize=4096,
                             isProvideTrainingMetric=True, metric="multi_logloss,multi_error")
    model = est.fit(LocalDataFrame.from_arrays(X, y, num_partitions=3))
    m = TextModel(model.model_str)
    out = model.transform(X[:8000])
    assert out["rawPrediction"].shape == (8000, 3)
    assert np.array_equal(out["rawPrediction"], m.predict_raw(X[:8000]))
    assert np.allclose(out["probability"], 1.0 / (1.0 + np.exp(-m.predict_raw(X[:8000]))), rtol=1e-12, atol=1e-15)
    assert (out["prediction"] == y[:8000]).mean() > 0.85
    last = est.performance_measures["evalHistory"][-1][0]
    p = m.predict(X)
    yi = y.astype(int)
    logloss = -np.log(np.maximum(p[np.arange(len(y)), yi], 1e-15)).mean()
    raw = m.predict_raw(X)
    err = ((raw >= raw[np.arange(len(y)), yi][:, None]).sum(axis=1) > 1).mean()
    assert abs(last["multi_logloss"] - logloss) < 1e-10 and abs(last["multi_error"] - err) < 1e-12
    # the model string reloads as a multiclassova model
    from synapseml_b200.lightgbm.booster import LightGBMBooster
    lb = LightGBMBooster(model_str=model.model_str)
    assert lb.num_classes() == 3
    assert np.allclose(lb.predict_for_mat(X[:100], LightGBMBooster.PREDICT_NORMAL).reshape(100, 3), out["probability"][:100], rtol=1e-12)
    lb.free_native_memory()
