"""Checks against implementations other than our own restatement.

* Wherever the real `lightgbm` Python package is importable, the oracle (CPU test) and the CUDA library (GPU test) are diffed against
  it: tree structure, thresholds (= bin upper bounds) and leaf values of the first trees.  The package is not installed in this image
  (no network), so these are skipped here and on the GPU box; they exist so that a maintainer with `pip install lightgbm==3.3.5` can pin
  the oracle — until then DESIGN.md says "parity unpinned".
* An accuracy sanity check (not parity): held-out AUC within 0.005 of scikit-learn's HistGradientBoostingClassifier, an independent
  histogram GBDT, with matching hyper-parameters on Higgs-shape data.
"""
import numpy as np
import pytest

from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.model_text import TextModel

PARAMS = dict(objective="binary", num_leaves=31, learning_rate=0.1, max_bin=255, min_data_in_leaf=20, min_sum_hessian_in_leaf=1e-3,
              boost_from_average=True, verbosity=-1, deterministic=True, force_col_wise=True, num_threads=1)


def _real_lightgbm_model(lgb, X, y, iters):
    ds = lgb.Dataset(X, label=y, params=dict(max_bin=255, bin_construct_sample_cnt=200000, min_data_in_bin=3, verbosity=-1), free_raw_data=False)
    bst = lgb.train(PARAMS, ds, num_boost_round=iters)
    return TextModel(bst.model_to_string())


def _assert_same_first_trees(ours, theirs, n_trees=3):
    so, st = ours.structure(), theirs.structure()
    assert so[:n_trees] == st[:n_trees]                 # split features, thresholds (bin upper bounds), topology
    for a, b in zip(ours.trees[:n_trees], theirs.trees[:n_trees]):
        assert np.allclose(a.leaf_value, b.leaf_value, rtol=1e-6, atol=1e-9)


def test_oracle_against_real_lightgbm_when_installed(oracle):
    lgb = pytest.importorskip("lightgbm")
    X, y = S.higgs_shape(50000)
    params = " ".join("%s=%s" % (k, str(v).lower() if isinstance(v, bool) else v) for k, v in PARAMS.items()
                      if k not in ("deterministic", "force_col_wise", "num_threads"))
    bst, dss = oracle.fit(X, y, params, num_iterations=5)
    _assert_same_first_trees(TextModel(bst.save_model_to_string()), _real_lightgbm_model(lgb, X, y, 5))


@pytest.mark.gpu
def test_library_against_real_lightgbm_when_installed(native):
    lgb = pytest.importorskip("lightgbm")
    from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LocalDataFrame
    X, y = S.higgs_shape(50000)
    est = LightGBMClassifier(numIterations=5, numLeaves=31, maxBin=255, minDataInLeaf=20, minSumHessianInLeaf=1e-3, samplingMode="fixed",
                             microBatchSize=8192)
    model = est.fit(LocalDataFrame.from_arrays(X, y))
    _assert_same_first_trees(TextModel(model.model_str), _real_lightgbm_model(lgb, X, y, 5))


@pytest.mark.gpu
@pytest.mark.parametrize("quantized", [False, True])
def test_auc_close_to_sklearn_hist_gradient_boosting(native, quantized):
    from sklearn.ensemble import HistGradientBoostingClassifier
    from sklearn.metrics import roc_auc_score
    from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LocalDataFrame
    X, y = S.higgs_shape(250000)
    Xtr, ytr, Xte, yte = X[:200000], y[:200000], X[200000:], y[200000:]
    est = LightGBMClassifier(numIterations=50, numLeaves=31, maxBin=255, learningRate=0.1, minDataInLeaf=20, samplingMode="fixed",
                             microBatchSize=1 << 15, useQuantizedGrad=quantized)
    ours = roc_auc_score(yte, est.fit(LocalDataFrame.from_arrays(Xtr, ytr)).predict_raw(Xte))
    sk = HistGradientBoostingClassifier(max_iter=50, max_leaf_nodes=31, learning_rate=0.1, max_bins=255, min_samples_leaf=20,
                                        early_stopping=False, random_state=0).fit(Xtr, ytr)
    theirs = roc_auc_score(yte, sk.decision_function(Xte))
    assert abs(ours - theirs) < 0.005, (ours, theirs)
