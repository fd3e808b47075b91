"""Categorical features (LightGBMBase.scala:162-198 resolves categoricalSlotIndexes / categoricalSlotNames to
the native categorical_feature list, LightGBMParams.scala:286-296,451-483) against the oracle's restatement of upstream's categorical
bin mapper (BinMapper::FindBin / ValueToBin, bin_type CategoricalBin), split search (FindBestThresholdCategoricalInner / ...IntInner:
one category against the rest for few categories, sorted many-against-many otherwise), bitset partition and Tree::SplitCategorical /
CategoricalDecision: bins bit-exact, quantized models byte-identical, scoring on raw values equal to an independent traversal."""
import ctypes as C
import json

import numpy as np
import pytest

from helpers import NativeBooster, NativeDataset, oracle_dataset, strip_parameters
from synapseml_b200.lightgbm.model_text import TextModel
from synapseml_b200.lightgbm.native import lightgbmlib, validate

pytestmark = pytest.mark.gpu


def cat_matrix(n=40000, seed=0):
    """Columns: 0 numeric, 1 three categories (one-hot search), 2 numeric with NaNs, 3 forty categories with NaN / negative values,
    4 a long-tailed categorical whose rare values fall into bin 0, 5 categories given as fractional values (cut to ints), 6 one
    category only (trivial), 7 numeric."""
    rng = np.random.Generator(np.random.PCG64(seed))
    X = rng.normal(size=(n, 8))
    X[:, 1] = rng.integers(0, 3, n)
    X[rng.random(n) < 0.1, 2] = np.nan
    X[:, 3] = rng.integers(0, 40, n)
    X[rng.random(n) < 0.05, 3] = np.nan
    X[rng.random(n) < 0.02, 3] = -3
    X[:, 4] = np.minimum(rng.geometric(0.02, n), 900) + 100           # hundreds of categories, most of them rare
    X[:, 5] = rng.integers(1, 12, n) + rng.random(n) * 0.9
    X[:, 6] = 7
    eff3, eff4, eff5 = rng.normal(size=40), rng.normal(size=1100), rng.normal(size=13)
    z = (X[:, 0] + eff3[np.nan_to_num(X[:, 3]).astype(int).clip(0)] + (X[:, 1] == 2) * 1.0 + 0.7 * eff4[X[:, 4].astype(int)]
         + 0.5 * eff5[X[:, 5].astype(int)] - 0.5 * np.nan_to_num(X[:, 2]))
    y = (z + rng.normal(0, 0.5, n) > 0).astype(np.float32)
    return X, y, z


CATS = "categorical_feature=1,3,4,5,6"


def test_categorical_bins_equal_the_oracle(native, oracle):
    X, y, _ = cat_matrix(30000)
    for extra in ("max_bin=255", "max_bin=15 min_data_in_bin=40", "max_bin=63 use_missing=false"):
        params = "objective=binary verbosity=-1 %s %s" % (CATS, extra)
        nd = NativeDataset(X, y, params, micro_batches=[4096, 33])
        ref, dss, _ = oracle_dataset(oracle, X, y, params)
        assert nd.num_used() == dss[0].num_used_feature()
        inner = 0
        for j in range(X.shape[1]):
            a, b = nd.feature_info(j), ref.feature_info(j)
            assert a == b, (extra, j, a, b)
            if a["is_trivial"]:
                continue
            assert np.array_equal(nd.bounds(j), ref.bin_upper_bounds(j), equal_nan=True), (extra, j)      # categorical: the category of every bin
            assert np.array_equal(nd.bin_column(inner), dss[0].bin_column(inner).astype(np.uint8)), (extra, j)
            inner += 1
        assert nd.feature_info(6)["is_trivial"]
        cats4 = nd.bounds(4)
        assert cats4[0] == -1 and len(cats4) < len(np.unique(X[:, 4]))       # the rare categories share bin 0
        nd.free()


def train_both(oracle, X, y, params, iters, weight=None):
    nd = NativeDataset(X, y, params, weight=weight, micro_batches=[8192])
    nb = NativeBooster(nd, params)
    _, dss, _ = oracle_dataset(oracle, X, y, params, weight=weight)
    ob = oracle.Booster(dss, params)
    for _ in range(iters):
        assert nb.update() == ob.update_one_iter()
    return nd, nb, ob


QUANT_CASES = {
    "default": "objective=binary num_leaves=31 max_bin=255 min_data_in_leaf=20",
    "small_groups": "objective=binary num_leaves=31 max_bin=63 min_data_per_group=10 cat_smooth=1 cat_l2=0.5 min_data_in_leaf=5",
    "few_categories_per_split": "objective=binary num_leaves=15 max_cat_threshold=3 max_bin=255",
    "all_onehot": "objective=binary num_leaves=31 max_cat_to_onehot=64 max_bin=63",
    "regularised": "objective=binary num_leaves=15 lambda_l1=0.5 lambda_l2=2.0 min_gain_to_split=0.1 max_delta_step=0.3 max_bin=63",
    "l2": "objective=regression num_leaves=31 max_bin=255 min_data_in_leaf=20",
    "stochastic": "objective=binary num_leaves=31 max_bin=255 stochastic_rounding=true seed=11",
    "column_sample_and_bagging": "objective=binary num_leaves=15 feature_fraction=0.6 bagging_freq=1 bagging_fraction=0.7 max_bin=63",
    "dart": "objective=binary num_leaves=15 boosting=dart drop_rate=0.3 skip_drop=0.2 max_bin=63",
    "multiclass": "objective=multiclass num_class=3 num_leaves=15 max_bin=63",
}


@pytest.mark.parametrize("case", sorted(QUANT_CASES))
def test_quantized_models_are_byte_identical(native, oracle, case):
    X, y, z = cat_matrix(40000)
    if case == "l2":
        y = z.astype(np.float32)
    if case == "multiclass":
        y = np.digitize(z, [-0.8, 0.8]).astype(np.float32)
    params = QUANT_CASES[case] + " use_quantized_grad=true verbosity=-1 " + CATS
    if "stochastic_rounding" not in params:
        params += " stochastic_rounding=false"
    nd, nb, ob = train_both(oracle, X, y, params, 10)
    ms = nb.model()
    assert strip_parameters(ms) == strip_parameters(ob.save_model_to_string())
    if case == "dart":
        assert np.allclose(nb.scores(), ob.scores(), rtol=0, atol=1e-12)
    else:
        assert np.array_equal(nb.scores(), ob.scores())
    m = TextModel(ms)
    assert sum(t.num_cat for t in m.trees if t.num_leaves > 1) > 0                 # categorical splits were found ...
    kinds = {int(d) & 1 for t in m.trees if t.num_leaves > 1 for d in t.decision_type}
    assert kinds == {0, 1}                                                         # ... next to numerical ones
    nb.free()
    nd.free()


def test_weights_and_float_mode(native, oracle):
    X, y, _ = cat_matrix(40000)
    rng = np.random.Generator(np.random.PCG64(4))
    w = rng.uniform(0.3, 2.0, len(y)).astype(np.float32)
    qp = "objective=binary num_leaves=31 max_bin=255 use_quantized_grad=true stochastic_rounding=false verbosity=-1 " + CATS
    nd, nb, ob = train_both(oracle, X, y, qp, 8, weight=w)
    assert strip_parameters(nb.model()) == strip_parameters(ob.save_model_to_string())
    nb.free(); nd.free()
    # float mode (19-bit fixed point against the oracle's double sums): the same first trees, scores within north_star's tolerance
    fp = "objective=binary num_leaves=15 max_bin=63 min_data_in_leaf=50 verbosity=-1 " + CATS
    nd, nb, ob = train_both(oracle, X, y, fp, 6)
    a, b = TextModel(nb.model()), TextModel(ob.save_model_to_string())
    assert a.structure()[:2] == b.structure()[:2]
    assert [t.cat_threshold.tolist() for t in a.trees[:2]] == [t.cat_threshold.tolist() for t in b.trees[:2]]
    assert np.allclose(nb.scores(), ob.scores(), rtol=1e-5, atol=1e-5)
    nb.free(); nd.free()


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


def test_scoring_contributions_reload_and_dump(native):
    from synapseml_b200.lightgbm.booster import LightGBMBooster
    X, y, _ = cat_matrix(30000)
    params = "objective=binary num_leaves=31 max_bin=255 learning_rate=0.2 verbosity=-1 " + CATS
    nd = NativeDataset(X, y, params, micro_batches=[4096])
    nb = NativeBooster(nd, params)
    for _ in range(10):
        nb.update()
    text = nb.model()
    tm = TextModel(text)
    raw = predict(nb.handle, X, 1)
    assert np.array_equal(raw, tm.predict_raw(X))                       # the same leaves, the same summation order
    assert np.allclose(raw, nb.scores(), rtol=0, atol=1e-12)            # the bin-level walk of training agrees with the value-level one
    # values no bin was made for: unseen, rare, negative, NaN and huge categories go right at every categorical node
    Xo = X[:2000].copy()
    Xo[:400, 3] = 77; Xo[400:800, 3] = -1; Xo[800:1200, 3] = np.nan; Xo[1200:1600, 3] = 1e12; Xo[1600:, 1] = 5
    assert np.array_equal(predict(nb.handle, Xo, 1), tm.predict_raw(Xo))
    leaves = predict(nb.handle, Xo, 2).reshape(len(Xo), -1)
    assert np.array_equal(leaves.astype(np.int64), np.stack([t.predict_leaf(Xo) for t in tm.trees], axis=1))
    # feature contributions walk the same decisions: they sum to the raw score
    contrib = predict(nb.handle, Xo[:500], 3).reshape(500, X.shape[1] + 1)
    assert np.allclose(contrib.sum(axis=1), predict(nb.handle, Xo[:500], 1), rtol=0, atol=1e-9)
    # model string round trip (LightGBMBooster.scala:41-49) and the JSON dump
    lb = LightGBMBooster(model_str=text)
    assert lb.save_to_string() == text
    assert np.array_equal(lb.predict_for_mat(Xo, LightGBMBooster.PREDICT_RAW), predict(nb.handle, Xo, 1))
    dump = json.loads(lb.dump_model())
    assert dump["feature_infos"]["Column_1"]["values"] == [int(v) for v in nd.bounds(1)]
    cat_nodes = []

    def walk(node):
        if "split_index" in node:
            if node["decision_type"] == "==":
                cat_nodes.append(node)
            walk(node["left_child"]); walk(node["right_child"])

    for t in dump["tree_info"]:
        assert t["num_cat"] == tm.trees[t["tree_index"]].num_cat
        walk(t["tree_structure"])
    assert cat_nodes and all(isinstance(nd_["threshold"], str) and nd_["threshold"] for nd_ in cat_nodes)
    first = cat_nodes[0]
    assert all(int(c) >= 0 for c in first["threshold"].split("||"))
    lb.free_native_memory()
    nb.free()
    nd.free()


def test_estimator_params_reach_the_native_side(native):
    from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LightGBMRegressor, LocalDataFrame
    X, y, z = cat_matrix(30000)
    names = ["c%d" % j for j in range(X.shape[1])]
    clf = LightGBMClassifier(numIterations=15, numLeaves=15, categoricalSlotIndexes=[1, 3], categoricalSlotNames=["c4", "c5"], slotNames=names,
                             samplingMode="fixed", microBatchSize=4096, maxCatThreshold=8, catSmooth=5.0, catl2=2.0, minDataPerGroup=50,
                             maxCatToOnehot=4)
    model = clf.fit(LocalDataFrame.from_arrays(X, y, num_partitions=3))
    text = model.getNativeModel()
    assert "[categorical_feature: 1,3,4,5]" in text and "[max_cat_threshold: 8]" in text and "[cat_smooth: 5" in text and "[cat_l2: 2" in text
    tm = TextModel(text)
    assert sum(t.num_cat for t in tm.trees if t.num_leaves > 1) > 0
    assert all(bin(int(w)).count("1") <= 8 for t in tm.trees if t.num_leaves > 1 for k in range(t.num_cat)
               for w in [sum(bin(int(x)).count("1") for x in t.cat_threshold[t.cat_boundaries[k]:t.cat_boundaries[k + 1]])])
    from sklearn.metrics import roc_auc_score
    assert roc_auc_score(y, model.predict_raw(X)) > 0.9
    # a name that is not a slot resolves to -1 (LightGBMBase.scala:171) and fails before training, as the reference's test expects
    bad = LightGBMRegressor(numIterations=2, categoricalSlotNames=["nope"], slotNames=names, samplingMode="fixed")
    with pytest.raises(Exception):
        bad.fit(LocalDataFrame.from_arrays(X, z.astype(np.float32)))


def test_bulk_and_sparse_ingest_with_categorical_and_wide_features(native):
    """dataTransferMode=bulk (LGBM_DatasetCreateFromMat, DatasetAggregator.scala:405-421) and sparse partitions (CSR push) build the
    same bins — and so the same model — as the dense streaming path, with categorical columns and features above 256 bins."""
    import scipy.sparse as sp
    from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LocalDataFrame
    X, y, _ = cat_matrix(30000)
    X = np.nan_to_num(X)                       # CSR rows cannot carry NaN-as-missing next to implicit zeros
    X[:, 3] = np.abs(X[:, 3])
    kw = dict(numIterations=6, numLeaves=15, maxBin=400, categoricalSlotIndexes=[1, 3, 5], useQuantizedGrad=True, stochasticRounding=False,
              samplingMode="fixed", microBatchSize=4096)
    stream = LightGBMClassifier(dataTransferMode="streaming", **kw).fit(LocalDataFrame.from_arrays(X, y, num_partitions=2))
    bulk = LightGBMClassifier(dataTransferMode="bulk", **kw).fit(LocalDataFrame.from_arrays(X, y, num_partitions=2))
    sparse = LightGBMClassifier(dataTransferMode="streaming", **kw).fit(
        LocalDataFrame([dict(features=sp.csr_matrix(X[:15000]), label=y[:15000]), dict(features=sp.csr_matrix(X[15000:]), label=y[15000:])]))
    a = strip_parameters(stream.getNativeModel())
    assert a == strip_parameters(bulk.getNativeModel()) == strip_parameters(sparse.getNativeModel())
    tm = TextModel(stream.getNativeModel())
    assert sum(t.num_cat for t in tm.trees if t.num_leaves > 1) > 0
