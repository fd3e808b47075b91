"""max_bin > 255 (maxBin is a free IntParam: LightGBMParams.scala:232-235; upstream then stores uint16 bins).  Here a feature with more
than 256 bins occupies several uint8 columns of the bin matrix (csrc/dataset.h), the row binning writes one coded byte per column and the
split search scans the columns as one feature; the histogram, partition and score kernels are the 8-bit ones.  Checked against the
oracle, which keeps such datasets as uint16: bins bit-exact, quantized models byte-identical, float mode within north_star's tolerance,
scoring on raw values equal to the training scores."""
import numpy as np
import pytest

from helpers import NativeBooster, NativeDataset, edge_case_matrix, oracle_dataset, strip_parameters
from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.model_text import TextModel

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("extra", ["max_bin=1023", "max_bin=511 zero_as_missing=true", "max_bin=300 min_data_in_bin=1", "max_bin=4000 use_missing=false"])
def test_wide_bins_equal_the_oracle(native, oracle, extra):
    X = edge_case_matrix(30000)           # continuous columns become wide, the few-valued ones stay ordinary 8-bit features
    y = (np.nan_to_num(X[:, 0]) > 0).astype(np.float32)
    params = "objective=binary verbosity=-1 " + extra
    nd = NativeDataset(X, y, params, micro_batches=[4096, 33])
    ref, dss, _ = oracle_dataset(oracle, X, y, params)
    assert nd.num_used() == dss[0].num_used_feature()
    inner, wide = 0, 0
    for j in range(X.shape[1]):
        a, b = nd.feature_info(j), ref.feature_info(j)
        assert a == b, (j, a, b)
        if a["is_trivial"]:
            continue
        assert np.array_equal(nd.bounds(j), ref.bin_upper_bounds(j), equal_nan=True), j
        assert np.array_equal(nd.bin_column(inner).astype(np.uint16), dss[0].bin_column(inner)), j
        wide += a["num_bin"] > 256
        inner += 1
    assert wide >= 3
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
    "max_bin_1023": "objective=binary num_leaves=31 max_bin=1023 min_data_in_leaf=20",
    "max_bin_511_deep": "objective=binary num_leaves=127 max_bin=511 min_data_in_leaf=5 learning_rate=0.2",
    "max_bin_300": "objective=binary num_leaves=31 max_bin=300",
    "regularised": "objective=binary num_leaves=15 lambda_l1=0.5 lambda_l2=2.0 min_gain_to_split=0.1 max_delta_step=0.3 max_bin=1023",
    "column_sample_and_bagging": "objective=binary num_leaves=15 feature_fraction=0.6 bagging_freq=1 bagging_fraction=0.7 max_bin=600",
    "stochastic": "objective=binary num_leaves=31 max_bin=1023 stochastic_rounding=true seed=3",
    "l2": "objective=regression num_leaves=31 max_bin=1023",
    "max_depth": "objective=binary num_leaves=31 max_depth=4 max_bin=2047",
}


@pytest.mark.parametrize("case", sorted(QUANT_CASES))
def test_quantized_models_are_byte_identical(native, oracle, case):
    if case == "l2":
        X, y = S.dense100(40000, f=20)
    else:
        X, y = S.higgs_shape(60000)      # continuous columns (wide) next to three-valued ones (ordinary)
    params = QUANT_CASES[case] + " use_quantized_grad=true verbosity=-1"
    if "stochastic_rounding" not in params:
        params += " stochastic_rounding=false"
    nd, nb, ob = train_both(oracle, X, y, params, 10)
    ms = nb.model()
    assert strip_parameters(ms) == strip_parameters(ob.save_model_to_string())
    assert np.array_equal(nb.scores(), ob.scores())
    assert max(nd.feature_info(j)["num_bin"] for j in range(X.shape[1])) > 256
    nb.free()
    nd.free()


@pytest.mark.parametrize("extra", ["", "zero_as_missing=true"])
def test_missing_values_weights_and_scoring(native, oracle, extra):
    X = edge_case_matrix(30000)
    rng = np.random.Generator(np.random.PCG64(2))
    z = np.nan_to_num(X[:, 0]) + np.isnan(X[:, 3]) * 1.5 - 0.5 * np.nan_to_num(X[:, 2]) + 0.3 * np.log1p(X[:, 8]) + (X[:, 1] == 0) * 0.7
    y = (rng.random(len(z)) < 1 / (1 + np.exp(-z))).astype(np.float32)
    w = rng.uniform(0.2, 3.0, len(z)).astype(np.float32)
    params = ("objective=binary num_leaves=31 max_bin=1023 min_data_in_leaf=20 use_quantized_grad=true stochastic_rounding=false verbosity=-1 "
              + extra)
    nd, nb, ob = train_both(oracle, X, y, params, 10, weight=w)
    ms = nb.model()
    assert strip_parameters(ms) == strip_parameters(ob.save_model_to_string())
    m = TextModel(ms)
    want = 1 if extra else 2
    assert any(((t.decision_type >> 2) & 3 == want).any() for t in m.trees if t.num_leaves > 1)      # splits on features with a missing bin
    assert np.allclose(m.predict_raw(X), nb.scores(), rtol=0, atol=1e-12)                            # the byte-coded walk agrees with the values
    nb.free()
    nd.free()


def test_float_mode_and_estimator(native, oracle):
    X, y = S.higgs_shape(60000)
    params = "objective=binary num_leaves=31 max_bin=1023 min_data_in_leaf=20 verbosity=-1"
    nd, nb, ob = train_both(oracle, X, y, params, 10)
    a, b = TextModel(nb.model()), TextModel(ob.save_model_to_string())
    assert a.structure()[:3] == b.structure()[:3]
    pa, pb = nb.scores(), ob.scores()
    assert np.abs(pa - pb).max() / np.abs(pb).max() < 1e-5
    nb.free()
    nd.free()
    from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LocalDataFrame
    from sklearn.metrics import roc_auc_score
    model = LightGBMClassifier(numIterations=20, numLeaves=31, maxBin=1023, samplingMode="fixed", microBatchSize=4096).fit(
        LocalDataFrame.from_arrays(X, y, num_partitions=2))
    assert "[max_bin: 1023]" in model.getNativeModel()
    narrow = LightGBMClassifier(numIterations=20, numLeaves=31, maxBin=255, samplingMode="fixed", microBatchSize=4096).fit(
        LocalDataFrame.from_arrays(X, y, num_partitions=2))
    assert abs(roc_auc_score(y, model.predict_raw(X)) - roc_auc_score(y, narrow.predict_raw(X))) < 0.01
