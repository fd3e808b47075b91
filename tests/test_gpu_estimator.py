"""Estimator API (the reference's user-facing surface) on the GPU: behaviours the reference's own suites assert."""
import numpy as np
import pytest

from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LightGBMRanker, LightGBMRegressor, LocalDataFrame

pytestmark = pytest.mark.gpu


def auc(y, p):
    order = np.argsort(p, kind="stable")
    r = np.empty(len(p))
    r[order] = np.arange(1, len(p) + 1)
    pos = y > 0
    return (r[pos].sum() - pos.sum() * (pos.sum() + 1) / 2) / (pos.sum() * (~pos).sum())


def test_classifier_fits_and_scores(native):
    X, y = S.higgs_shape(60000)
    df = LocalDataFrame.from_arrays(X, y, num_partitions=2)            # numPartitions = 2, LightGBMTestUtils.scala:95
    model = LightGBMClassifier(numIterations=10, numLeaves=5).fit(df)   # the reference's benchmark setting (10 iters, 5 leaves)
    out = model.transform(X)
    assert auc(y, out["probability"][:, 1]) > 0.72
    assert out["rawPrediction"].shape == (len(y), 2) and np.allclose(out["rawPrediction"][:, 0], -out["rawPrediction"][:, 1])
    assert "is_unbalance: false" in model.getNativeModel() and "[learning_rate: 0.1]" in model.getNativeModel()


def test_streaming_and_bulk_give_the_same_model(native):
    # the reference re-runs its streaming suites in bulk mode by subclass override (SURVEY.md §4 "mode matrix")
    X, y = S.higgs_shape(30000)
    df = LocalDataFrame.from_arrays(X, y, num_partitions=3)
    kw = dict(numIterations=5, numLeaves=15, useQuantizedGrad=True, stochasticRounding=False, samplingMode="fixed")
    a = LightGBMClassifier(dataTransferMode="streaming", **kw).fit(df).getNativeModel().split("parameters:")[0]
    b = LightGBMClassifier(dataTransferMode="bulk", **kw).fit(df).getNativeModel().split("parameters:")[0]
    assert a == b


def test_micro_batch_size_invariance_and_threaded_push(native):
    # VerifyLightGBMClassifierStream.scala:395-405 (chunk size invariance)
    X, y = S.higgs_shape(20000)
    df = LocalDataFrame.from_arrays(X, y, num_partitions=4)
    kw = dict(numIterations=4, numLeaves=15, useQuantizedGrad=True, stochasticRounding=False, samplingMode="fixed")
    models = [LightGBMClassifier(microBatchSize=mb, **kw).fit(df).getNativeModel() for mb in (1, 100, 4096)]
    assert models[0] == models[1] == models[2]


def test_reference_dataset_reuse_gives_identical_model_string(native):
    # VerifyLightGBMClassifierStreamOnly.scala:44-56
    X, y = S.higgs_shape(20000)
    df = LocalDataFrame.from_arrays(X, y, num_partitions=2)
    est = LightGBMClassifier(numIterations=4, numLeaves=15, samplingMode="fixed")
    m1 = est.fit(df)
    assert est.getReferenceDataset() is not None
    est2 = LightGBMClassifier(numIterations=4, numLeaves=15, samplingMode="fixed", referenceDataset=est.getReferenceDataset())
    assert est2.fit(df).getNativeModel() == m1.getNativeModel()


def test_determinism(native):
    # VerifyLightGBMClassifierStream.scala:95-101 (same input -> same result, 1e-10); here: identical model strings,
    # float mode included, because histograms are integer sums in every mode
    X, y = S.higgs_shape(50000)
    df = LocalDataFrame.from_arrays(X, y, num_partitions=2)
    ms = [LightGBMClassifier(numIterations=6, seed=1, deterministic=True).fit(df).getNativeModel() for _ in range(3)]
    assert ms[0] == ms[1] == ms[2]


def test_empty_partition_and_sampling_modes(native):
    # VerifyLightGBMClassifierStream.scala:449-461 and VerifyLightGBMClassifierStreamOnly.scala:12-42
    X, y = S.higgs_shape(12000)
    parts = LocalDataFrame.from_arrays(X, y, num_partitions=2).partitions
    parts.insert(1, {"features": X[:0], "label": y[:0]})
    for mode in ("global", "fixed", "subset"):
        m = LightGBMClassifier(numIterations=3, numLeaves=7, samplingMode=mode).fit(LocalDataFrame(parts))
        assert len(m._model.trees) == 3


def test_regressor_and_weight_column(native):
    X, y = S.dense100(30000, f=20)
    df = LocalDataFrame.from_arrays(X, y, num_partitions=2)
    m = LightGBMRegressor(numIterations=20, numLeaves=15).fit(df)
    rmse0 = np.sqrt(np.mean((y - y.mean()) ** 2))
    rmse = np.sqrt(np.mean((y - m.transform(X)["prediction"]) ** 2))
    assert rmse < 0.6 * rmse0
    # a weight column that favours large labels shifts the mean prediction up (VerifyLightGBMRegressorStream "weight column")
    w = np.where(y > np.median(y), 10.0, 0.1).astype(np.float32)
    mw = LightGBMRegressor(numIterations=20, numLeaves=15).fit(LocalDataFrame.from_arrays(X, y, num_partitions=2, weight=w))
    assert mw.transform(X)["prediction"].mean() > m.transform(X)["prediction"].mean()


def test_ranker_group_column_types(native):
    # VerifyLightGBMRankerStream.scala:22-74: group column as int, long and string; rejects a double group column
    X, rel, qid = S.mslr_shape(12000, f=20)
    ms = []
    for conv in (lambda q: q.astype(np.int32), lambda q: q.astype(np.int64) + (1 << 40), lambda q: np.array(["q%06d" % v for v in q])):
        df = LocalDataFrame.from_arrays(X, rel, num_partitions=1, group=conv(qid))
        ms.append(LightGBMRanker(numIterations=4, numLeaves=15, minDataInLeaf=10).fit(df).getNativeModel())
    assert ms[0] == ms[1] == ms[2]
    with pytest.raises(ValueError):
        LightGBMRanker(numIterations=1).fit(LocalDataFrame.from_arrays(X, rel, num_partitions=1, group=qid.astype(np.float64)))


def test_ranker_with_rank_xendcg_and_objective_seed(native):
    """objective=rank_xendcg with objectiveSeed (LightGBMParams.scala:441-445 exists for this objective alone): the per-query random
    streams make the model depend on the seed, and reproducible for a given one."""
    X, rel, qid = S.mslr_shape(12000, f=20)
    df = LocalDataFrame.from_arrays(X, rel, num_partitions=1, group=qid.astype(np.int32))
    kw = dict(objective="rank_xendcg", numIterations=6, numLeaves=15, minDataInLeaf=10, evalAt=[1, 3], isProvideTrainingMetric=True)
    est = LightGBMRanker(objectiveSeed=7, **kw)
    a = est.fit(df).getNativeModel()
    assert "objective=rank_xendcg" in a and "[objective_seed: 7]" in a
    assert a == LightGBMRanker(objectiveSeed=7, **kw).fit(df).getNativeModel()
    assert a.split("parameters:")[0] != LightGBMRanker(objectiveSeed=8, **kw).fit(df).getNativeModel().split("parameters:")[0]
    hist = est.performance_measures["evalHistory"]
    assert hist[-1][0]["ndcg@3"] > hist[0][0]["ndcg@3"] > 0.3          # the default metric of the objective is ndcg, and it improves


def test_unsupported_options_fail_loudly(native):
    X, y = S.higgs_shape(2000)
    df = LocalDataFrame.from_arrays(X, y)
    from synapseml_b200.lightgbm.native import LightGBMNativeError
    for kw in (dict(boostingType="plain"), dict(objective="cross_entropy_lambda"), dict(maxBin=100000), dict(passThroughArgs="feature_fraction_bynode=1.5"),
               dict(passThroughArgs="monotone_constraints=1,0,0")):
        with pytest.raises((LightGBMNativeError, NotImplementedError)):
            LightGBMClassifier(numIterations=1, **kw).fit(df)


def test_bagging_and_feature_fraction_params_reach_the_learner(native):
    # VerifyLightGBMClassifierStream.scala "bagging" / featureFraction settings: they change the model and are echoed in it
    X, y = S.higgs_shape(40000)
    df = LocalDataFrame.from_arrays(X, y, num_partitions=2)
    kw = dict(numIterations=8, numLeaves=15, useQuantizedGrad=True, stochasticRounding=False, samplingMode="fixed")
    plain = LightGBMClassifier(**kw).fit(df)
    sampled = LightGBMClassifier(baggingFraction=0.7, baggingFreq=1, baggingSeed=3, featureFraction=0.6, **kw).fit(df)
    again = LightGBMClassifier(baggingFraction=0.7, baggingFreq=1, baggingSeed=3, featureFraction=0.6, **kw).fit(df)
    ms = sampled.getNativeModel()
    assert "[bagging_fraction: 0.7]" in ms and "[feature_fraction: 0.6]" in ms and "[bagging_freq: 1]" in ms
    assert ms == again.getNativeModel()
    assert ms.split("parameters:")[0] != plain.getNativeModel().split("parameters:")[0]
    assert auc(y, sampled.transform(X)["probability"][:, 1]) > 0.72


def test_sparse_partitions_give_the_dense_model(native):
    # matrixType sparse: CSR micro-batches (StreamingPartitionTask.scala:239-277) must bin exactly like the dense rows
    import scipy.sparse as sp
    X, y = S.higgs_shape(20000)
    X = np.where(np.abs(X) < 0.7, 0.0, X)           # ~half zeros
    kw = dict(numIterations=4, numLeaves=15, useQuantizedGrad=True, stochasticRounding=False, samplingMode="fixed", microBatchSize=777)
    dense = LightGBMClassifier(**kw).fit(LocalDataFrame.from_arrays(X, y, num_partitions=3)).getNativeModel()
    parts = LocalDataFrame.from_arrays(X, y, num_partitions=3).partitions
    for p in parts:
        p["features"] = sp.csr_matrix(p["features"])
    sparse = LightGBMClassifier(**kw).fit(LocalDataFrame(parts)).getNativeModel()
    assert sparse == dense
    bulk_sparse = LightGBMClassifier(dataTransferMode="bulk", **kw).fit(LocalDataFrame(parts)).getNativeModel()
    assert bulk_sparse.split("parameters:")[0] == dense.split("parameters:")[0]



def test_model_methods(native, tmp_path):
    """LightGBMModelMethods.scala:19-107 and the models' companion objects (loadNativeModelFromFile / ...FromString)."""
    from synapseml_b200.lightgbm.estimators import LightGBMClassificationModel
    X, y = S.higgs_shape(20000)
    model = LightGBMClassifier(numIterations=7, numLeaves=15, samplingMode="fixed").fit(LocalDataFrame.from_arrays(X, y))
    assert model.getBoosterNumTotalIterations() == 7 and model.getBoosterNumTotalModel() == 7
    assert model.getBoosterNumFeatures() == X.shape[1] and model.getBoosterNumClasses() == 1
    assert model.getBoosterBestIteration() == -1                          # no early stopping
    shap = np.array(model.getFeatureShaps(X[3]))
    assert shap.shape == (X.shape[1] + 1,) and np.isclose(shap.sum(), model.predict_raw(X[3:4])[0], atol=1e-9)
    nz = np.flatnonzero(X[3])
    assert np.allclose(model.getSparseFeatureShaps(X.shape[1], nz, X[3][nz]), shap) and np.allclose(model.getDenseFeatureShaps(X[3]), shap)
    imp = model.getFeatureImportances("split")
    assert len(imp) == X.shape[1] and sum(imp) == sum(t.num_leaves - 1 for t in model._model.trees)
    path = tmp_path / "model.txt"
    model.saveNativeModel(str(path))
    for again in (LightGBMClassificationModel.loadNativeModelFromFile(str(path)), LightGBMClassificationModel.loadNativeModelFromString(model.getNativeModel())):
        assert again.getNativeModel() == model.getNativeModel()
        assert np.array_equal(again.transform(X[:500])["probability"], model.transform(X[:500])["probability"])


def test_dart_seed_batch_and_prediction_params(native):
    """LightGBMDartParams, the remaining LightGBMSeedParams, numBatches (LightGBMBase.scala:39-64) and LightGBMPredictionParams
    (leafPredictionCol / featuresShapCol / startIteration)."""
    X, y = S.higgs_shape(40000)
    df = LocalDataFrame.from_arrays(X, y, num_partitions=2)
    kw = dict(numIterations=10, numLeaves=15, samplingMode="fixed")
    dart = LightGBMClassifier(boostingType="dart", dropRate=0.4, maxDrop=3, skipDrop=0.1, uniformDrop=True, xGBoostDartMode=True, dropSeed=9,
                              featureFractionSeed=5, featureFraction=0.8, leafPredictionCol="leaves", featuresShapCol="shap", **kw)
    model = dart.fit(df)
    text = model.getNativeModel()
    for frag in ("[drop_rate: 0.4]", "[max_drop: 3]", "[skip_drop: 0.1]", "[uniform_drop: true]", "[xgboost_dart_mode: true]", "[drop_seed: 9]",
                 "[feature_fraction_seed: 5]"):
        assert frag in text, frag
    out = model.transform(X[:300])
    assert out["leaves"].shape == (300, 10) and out["shap"].shape == (300, X.shape[1] + 1)
    assert np.allclose(out["shap"].sum(axis=1), out["rawPrediction"][:, 1], atol=1e-9)
    # startIteration: the first trees are left out of the prediction
    late = model.setStartIteration(4).transform(X[:300])
    assert late["leaves"].shape == (300, 6) and not np.allclose(late["rawPrediction"], out["rawPrediction"])
    # numBatches: three models trained in sequence, each continuing from the one before: 3 x 10 iterations in the final model
    batched = LightGBMClassifier(numBatches=3, seed=4, **kw)
    bm = batched.fit(df)
    assert bm.getBoosterNumTotalIterations() == 30 and len(batched.batch_performance_measures) == 3
    from sklearn.metrics import roc_auc_score
    assert roc_auc_score(y, bm.predict_raw(X)) > roc_auc_score(y, LightGBMClassifier(**kw).fit(df).predict_raw(X)) - 0.01
    with pytest.raises(ValueError):
        LightGBMClassifier(objective="multiclass", maxNumClasses=1, **kw).fit(df)


def test_predict_disable_shape_check(native):
    """predictDisableShapeCheck (LightGBMParams.scala:344-349 -> predict_disable_shape_check in the scoring parameter string,
    LightGBMBooster.scala:526): rows wider or narrower than the training data are refused unless the check is disabled; then extra
    columns are ignored and missing ones read as 0."""
    X, y = S.higgs_shape(20000)
    model = LightGBMClassifier(numIterations=5, numLeaves=15, samplingMode="fixed").fit(LocalDataFrame.from_arrays(X, y))
    want = model.predict_raw(X[:200])
    wide = np.concatenate([X[:200], np.ones((200, 3))], axis=1)
    from synapseml_b200.lightgbm.native import LightGBMNativeError
    with pytest.raises(LightGBMNativeError):
        model.predict_raw(wide)
    model.setPredictDisableShapeCheck(True)
    assert np.array_equal(model.predict_raw(wide), want)
    narrow = X[:200, :20]
    padded = np.concatenate([narrow, np.zeros((200, 8))], axis=1)
    assert np.array_equal(model.predict_raw(narrow), model.predict_raw(padded))


def test_prediction_columns_and_thresholds(native):
    # Spark's predictor columns on the classification model (LightGBMClassifier.scala:118-163): renamed columns, a column left out by
    # an empty name, thresholds re-weighting the predicted class (ProbabilisticClassificationModel.probability2prediction)
    X, y = S.higgs_shape(20000)
    df = LocalDataFrame.from_arrays(X, y, num_partitions=1)
    model = LightGBMClassifier(numIterations=8, numLeaves=15, predictionCol="pred", rawPredictionCol="", probabilityCol="prob").fit(df)
    out = model.transform(X[:4000])
    assert set(out) == {"pred", "prob"}
    p = out["prob"][:, 1]
    assert np.array_equal(out["pred"], (p > 0.5).astype(np.float64))
    assert model.getActualNumClasses() == 2 and model.numClasses == 2 and model.getLightGBMBooster().num_classes() == 1
    model.setThresholds([0.8, 0.2])
    shifted = model.transform(X[:4000])["pred"]
    assert np.array_equal(shifted, (p / 0.2 > (1 - p) / 0.8).astype(np.float64)) and shifted.sum() > out["pred"].sum()
    model.setThresholds([0.0, 1.0])       # a zero threshold always wins (its probability is never exactly 0 here)
    assert not model.transform(X[:4000])["pred"].any()
    with pytest.raises(ValueError):
        model.setThresholds([0.5, 0.3, 0.2]).transform(X[:10])
    # multiclass, thresholds through the estimator
    rng = np.random.Generator(np.random.PCG64(3))
    y3 = np.digitize(X[:, 0] + 0.3 * rng.normal(size=len(X)), [-0.5, 0.6]).astype(np.float64)
    m3 = LightGBMClassifier(objective="multiclass", numIterations=5, numLeaves=7, thresholds=[1.0, 0.1, 1.0]).fit(LocalDataFrame.from_arrays(X, y3))
    o3 = m3.transform(X[:3000])
    assert m3.getActualNumClasses() == 3
    assert np.array_equal(o3["prediction"], np.argmax(o3["probability"] / np.array([1.0, 0.1, 1.0]), axis=1).astype(np.float64))
    assert (o3["prediction"] == 1).mean() > (np.argmax(o3["rawPrediction"], axis=1) == 1).mean()
    # the regressor's only column
    r = LightGBMRegressor(numIterations=3, numLeaves=7, predictionCol="yhat").fit(df).transform(X[:100])
    assert set(r) == {"yhat"}


def test_booster_wrapper_methods_and_prediction_window(native, tmp_path):
    # LightGBMBooster.scala:394-518: score's class expansion, predictLeaf / featuresShap of one row, the start / count window
    # setStartIteration / setNumIterations / setBestIteration put on every prediction, saveNativeModel, the iteration counts
    from synapseml_b200.lightgbm.booster import LightGBMBooster
    X, y = S.higgs_shape(20000)
    model = LightGBMClassifier(numIterations=6, numLeaves=7).fit(LocalDataFrame.from_arrays(X, y))
    b = model.getLightGBMBooster()
    assert b.num_total_iterations() == 6 and b.num_model_per_iteration() == 1 and b.get_native_model() == model.getNativeModel()
    r, p = b.score(X[5], raw=True), b.score(X[5])
    assert b.score(X[5], raw=True, classification=True) == [-r, r] and b.score(X[5], classification=True) == [1 - p, p]
    assert b.score(X[5], raw=True, classification=False) == [r]
    leaves = b.predict_leaf(X[5])
    assert len(leaves) == 6 and leaves == model.predict_leaf(X[5:6])[0].tolist()
    shap = b.features_shap(X[5])
    assert len(shap) == X.shape[1] + 1 and abs(sum(shap) - r) < 1e-9
    b.save_native_model(str(tmp_path / "m.txt"))
    assert open(tmp_path / "m.txt").read() == model.getNativeModel()
    with pytest.raises(FileExistsError):
        b.save_native_model(str(tmp_path / "m.txt"), overwrite=False)
    with pytest.raises(ValueError):
        b.save_native_model("")
    # the window: the model's numIterations / startIteration reach the native predictor
    full = model.predict_raw(X[:500])
    first2 = model.setNumIterations(2).predict_raw(X[:500])
    rest = model.setNumIterations(-1).setStartIteration(2).predict_raw(X[:500])
    model.setStartIteration(0)
    two = LightGBMBooster(model_str=b.save_to_string(2))
    assert np.array_equal(first2, two.predict_for_mat(X[:500], LightGBMBooster.PREDICT_RAW))
    assert np.allclose(first2 + rest, full, rtol=0, atol=1e-12) and not np.array_equal(first2, full)
    two.set_best_iteration(1)
    assert two.best_iteration == 1 and len(two.predict_leaf(X[0])) == 1
    two.free_native_memory()


def test_early_stopped_model_predicts_with_best_iteration_trees(native):
    # getModel(...).setNumIterations(booster.bestIteration) (LightGBMClassifier.scala:80): the model text keeps bestIteration + 1
    # iterations (saveToString, LightGBMBooster.scala:272-278) and transform uses the first bestIteration of them
    X, y = S.higgs_shape(30000)
    rng = np.random.Generator(np.random.PCG64(9))
    y = np.where(rng.random(len(y)) < 0.35, 1 - y, y).astype(np.float32)
    valid = np.zeros(len(y), dtype=bool)
    valid[20000:] = True
    est = LightGBMClassifier(numIterations=300, numLeaves=15, learningRate=0.1, validationIndicatorCol="isVal",
                             earlyStoppingRound=3, metric="binary_logloss", samplingMode="fixed")
    model = est.fit(LocalDataFrame.from_arrays(X, y, num_partitions=2, isVal=valid))
    best = est.best_iteration
    assert best is not None and best >= 1 and model.getBoosterBestIteration() == best and model.numIterations == best
    assert model.getBoosterNumTotalIterations() == best + 1
    from synapseml_b200.lightgbm.booster import LightGBMBooster
    lb = LightGBMBooster(model_str=model.getNativeModel())
    assert np.array_equal(model.predict_raw(X[:300]), lb.predict_for_mat(X[:300], LightGBMBooster.PREDICT_RAW, num_iteration=best))
    assert model.predict_leaf(X[:5]).shape == (5, best)
    lb.free_native_memory()
