"""Continued training (TrainUtils.scala:22-24, LightGBMBooster.scala:254-258): LGBM_BoosterMerge puts the
trees of an earlier model in front of the booster's own, as upstream's GBDT::MergeFrom does (scores are not touched, so the new
trees are the ones a fresh booster without an automatic initial score grows)."""
import ctypes as C

import numpy as np
import pytest

from helpers import NativeBooster, NativeDataset
from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LocalDataFrame
from synapseml_b200.lightgbm.model_text import TextModel
from synapseml_b200.lightgbm.native import lightgbmlib, validate

pytestmark = pytest.mark.gpu
PARAMS = ("objective=binary num_leaves=15 max_bin=63 min_data_in_leaf=20 use_quantized_grad=true stochastic_rounding=false "
          "verbosity=-1")


def test_merged_trees_lead_and_new_trees_follow(native):
    X, y = S.higgs_shape(30000)
    nd = NativeDataset(X, y, PARAMS)
    a = NativeBooster(nd, PARAMS)
    for _ in range(5):
        a.update()
    first = a.model()
    lib = lightgbmlib()
    loaded, n_iter = C.c_void_p(), C.c_int()
    validate(lib.LGBM_BoosterLoadModelFromString(first.encode(), C.byref(n_iter), C.byref(loaded)))
    b = NativeBooster(nd, PARAMS)
    validate(lib.LGBM_BoosterMerge(b.handle, loaded))
    validate(lib.LGBM_BoosterFree(loaded))
    cur = C.c_int()
    validate(lib.LGBM_BoosterGetCurrentIteration(b.handle, C.byref(cur)))
    assert cur.value == 5
    for _ in range(3):
        b.update()
    fresh = NativeBooster(nd, PARAMS + " boost_from_average=false")
    for _ in range(3):
        fresh.update()
    ma, mb, mf = TextModel(first), TextModel(b.model()), TextModel(fresh.model())
    assert len(mb.trees) == 8
    assert mb.structure()[:5] == ma.structure() and mb.structure()[5:] == mf.structure()
    for t, u in zip(mb.trees[:5] + mb.trees[5:], ma.trees + mf.trees):
        assert np.array_equal(t.leaf_value, u.leaf_value)
    assert np.array_equal(b.scores(), fresh.scores())          # the merge adds nothing to the training scores
    assert np.allclose(mb.predict_raw(X), ma.predict_raw(X) + mf.predict_raw(X), rtol=0, atol=1e-12)
    for nb in (a, b, fresh):
        nb.free()
    nd.free()


def test_estimator_model_string_continues_training(native):
    X, y = S.higgs_shape(20000)
    df = LocalDataFrame.from_arrays(X, y, num_partitions=1)
    kw = dict(numLeaves=15, useQuantizedGrad=True, stochasticRounding=False, samplingMode="fixed")
    m1 = LightGBMClassifier(numIterations=4, **kw).fit(df)
    m2 = LightGBMClassifier(numIterations=3, modelString=m1.getNativeModel(), **kw).fit(df)
    t1, t2 = TextModel(m1.getNativeModel()), TextModel(m2.getNativeModel())
    assert len(t2.trees) == 7 and t2.structure()[:4] == t1.structure()
