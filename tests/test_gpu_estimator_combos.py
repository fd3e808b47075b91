"""The Python mirror of the Scala layer against the bare C ABI: seeded random estimator configurations — estimator kind, data
transfer mode, dense / sparse partitions, partition count, micro-batch size, weight / init-score / group columns, categorical slots,
bin width, boosting type and the learner's knobs — are fitted through LightGBMClassifier / Regressor / Ranker and, independently,
through the LGBM_* call sequence on the same rows with the parameter strings the estimator builds.  The two models must be the same
text: the mirror adds nothing and loses nothing on the way (param plumbing, sampling, partition tasks, metadata columns)."""
import numpy as np
import pytest
import scipy.sparse as sp

from helpers import NativeBooster, NativeDataset, strip_parameters
from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LightGBMRanker, LightGBMRegressor, LocalDataFrame
from synapseml_b200.lightgbm.params import dataset_creation_params

pytestmark = pytest.mark.gpu


def make_data(n, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    X = rng.normal(size=(n, 9))
    X[:, 1] = rng.integers(0, 5, n)
    X[:, 2] = np.where(rng.random(n) < 0.6, 0.0, rng.exponential(1.0, n))
    X[:, 3] = rng.integers(0, 25, n)
    X[:, 4] = np.where(rng.random(n) < 0.5, 0.0, X[:, 4])
    eff = rng.normal(size=25)
    z = X[:, 0] - 0.6 * X[:, 5] + eff[X[:, 3].astype(int)] + (X[:, 1] == 3) * 0.9 + 0.5 * X[:, 2]
    return X, z, rng


@pytest.mark.parametrize("seed", range(40))
def test_estimator_fit_equals_the_bare_call_sequence(native, seed):
    r = np.random.Generator(np.random.PCG64(9000 + seed))
    pick = lambda *xs: xs[int(r.integers(0, len(xs)))]
    n = int(pick(6000, 9000))
    X, z, rng = make_data(n, seed)
    kind = pick("classifier", "classifier", "regressor", "ranker")
    kw = dict(numIterations=int(pick(3, 5)), numLeaves=int(pick(7, 15, 31)), maxBin=int(pick(15, 63, 255, 400)), minDataInLeaf=int(pick(5, 20)),
              learningRate=float(pick(0.1, 0.3)), samplingMode="fixed", microBatchSize=int(pick(1, 100, 5000)),
              dataTransferMode=pick("streaming", "streaming", "bulk"), useQuantizedGrad=True, stochasticRounding=False,
              useSingleDatasetMode=bool(pick(True, False)), chunkSize=int(pick(1000, 10000)))
    if pick(True, False):
        kw["categoricalSlotIndexes"] = pick([1, 3], [3], [1])
    if pick(True, False, False):
        kw.update(baggingFraction=0.7, baggingFreq=1, baggingSeed=int(pick(3, 9)))
    if pick(True, False, False):
        kw.update(featureFraction=0.7)
    if pick(True, False, False):
        kw.update(lambdaL1=0.2, lambdaL2=1.0)
    if pick(True, False, False):
        kw["maxDepth"] = 4
    if pick(True, False, False, False):
        kw["zeroAsMissing"] = True
    boosting = pick("gbdt", "gbdt", "gbdt", "dart", "goss", "rf")
    kw["boostingType"] = boosting
    if boosting == "rf":
        kw.update(baggingFraction=0.7, baggingFreq=1, featureFraction=0.8)
    if boosting == "goss":
        for k in ("baggingFraction", "baggingFreq", "baggingSeed"):
            kw.pop(k, None)
    if pick(True, False, False):
        kw["seed"] = int(pick(1, 7, 42))
    weight = rng.uniform(0.5, 2.0, n).astype(np.float64) if pick(True, False, False) else None
    init = (0.3 * rng.normal(size=n)) if pick(True, False, False, False) else None
    cols, group = {}, None
    if kind == "classifier":
        objective = pick("binary", "binary", "multiclass", "multiclassova")
        y = (z + rng.normal(0, 0.5, n) > 0).astype(np.float64) if objective == "binary" else np.digitize(z + rng.normal(0, 0.3, n), [-0.8, 0.9]).astype(np.float64)
        if objective != "binary":
            init = None          # (class-major init scores are covered by test_gpu_multiclass)
        est = LightGBMClassifier(objective=objective, **kw)
    elif kind == "regressor":
        objective = pick("regression", "regression_l1", "huber", "quantile", "tweedie")
        y = np.abs(z) + 0.1 if objective == "tweedie" else z + rng.normal(0, 0.3, n)
        est = LightGBMRegressor(objective=objective, **kw)
    else:
        objective = pick("lambdarank", "rank_xendcg")
        y = np.digitize(z + rng.normal(0, 0.5, n), [-0.5, 0.8, 2.0]).astype(np.float64)
        sizes = []
        while sum(sizes) < n:
            sizes.append(int(rng.integers(3, 40)))
        group = np.repeat(np.arange(len(sizes)), sizes)[:n].astype(np.int32)
        cols["group"] = group
        init = None
        est = LightGBMRanker(objective=objective, **kw)
    if weight is not None:
        cols["weight"] = weight
        est._set("weightCol", "weight")
    if init is not None:
        cols["init"] = init
        est._set("initScoreCol", "init")
    sparse = kw["dataTransferMode"] == "streaming" and "zeroAsMissing" not in kw and pick(True, False, False)
    nparts = 1 if kind == "ranker" else int(pick(1, 2, 3))
    if sparse:
        bounds = [n * i // nparts for i in range(nparts + 1)]
        parts = []
        for b, e in zip(bounds[:-1], bounds[1:]):
            p = {"features": sp.csr_matrix(X[b:e]), "label": y[b:e]}
            p.update({k: v[b:e] for k, v in cols.items()})
            parts.append(p)
        df = LocalDataFrame(parts)
    else:
        df = LocalDataFrame.from_arrays(X, y, num_partitions=nparts, **cols)
    model = est.fit(df)

    # the same fit through the bare C ABI with the estimator's own parameter strings
    g = est._p
    tp = est.add_custom_train_params(est.get_train_params(1, max(1, nparts)), df, None)
    ds_params = dataset_creation_params(est.getMaxBin(), est.getBinSampleCount(), est.getMinDataInLeaf(), tp.executionParams.numThreads,
                                        categorical_indexes=est.get_categorical_indexes(),
                                        data_random_seed=g["dataRandomSeed"] if g["dataRandomSeed"] is not None else g["seed"])
    nd = NativeDataset(X, y, ds_params, weight=None if weight is None else weight.astype(np.float32), query=group,
                       init_score=init, micro_batches=[4096])
    nb = NativeBooster(nd, tp.to_string())
    for _ in range(kw["numIterations"]):
        if nb.update():
            break
    assert strip_parameters(model.getNativeModel()) == strip_parameters(nb.model()), (kind, objective, kw)
    nb.free()
    nd.free()
