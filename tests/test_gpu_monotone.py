"""Monotone constraints (monotoneConstraints / monotoneConstraintsMethod / monotonePenalty: LightGBMParams.scala, GeneralParams in
params/TrainParams.scala), method "basic": upstream BasicLeafConstraints + the USE_MC split search (gains at outputs clamped to the
leaf's range, direction check, children's ranges cut at the mean of the two outputs, depth penalty).  Against the oracle: quantized
models byte-identical; and against the property itself: predictions are monotone in the constrained features."""
import numpy as np
import pytest

from helpers import NativeBooster, NativeDataset, oracle_dataset, strip_parameters
from synapseml_b200.lightgbm.model_text import TextModel

pytestmark = pytest.mark.gpu


def mono_data(n=40000, seed=0):
    rng = np.random.Generator(np.random.PCG64(seed))
    X = rng.normal(size=(n, 6))
    X[:, 4] = rng.integers(0, 12, n)
    X[rng.random(n) < 0.1, 1] = np.nan
    eff = rng.normal(size=12)
    z = X[:, 0] + 0.8 * np.sin(3 * X[:, 0]) - np.nan_to_num(X[:, 1]) + 0.3 * X[:, 2] ** 2 + 0.5 * eff[X[:, 4].astype(int)]
    return X, z, rng


def train_both(oracle, X, y, params, iters):
    nd = NativeDataset(X, y, params, micro_batches=[8192])
    nb = NativeBooster(nd, params)
    _, dss, _ = oracle_dataset(oracle, X, y, params)
    ob = oracle.Booster(dss, params)
    for _ in range(iters):
        assert nb.update() == ob.update_one_iter()
    return nd, nb, ob


def assert_monotone(tm, X, col, sign):
    base = X[:300].copy()
    grid = np.linspace(-3, 3, 50)
    prev = None
    for g in grid:
        a = base.copy()
        a[:, col] = g
        p = tm.predict_raw(a)
        if prev is not None:
            assert (sign * (p - prev) >= -1e-12).all()
        prev = p


CASES = {
    "regression": "objective=regression num_leaves=31 max_bin=255 min_data_in_leaf=20 monotone_constraints=1,-1,0,0,0,0",
    "penalty": "objective=regression num_leaves=31 max_bin=63 monotone_constraints=1,-1,0,0,0,0 monotone_penalty=1.5",
    "small_penalty": "objective=regression num_leaves=15 max_bin=63 monotone_constraints=1,-1,0,0,0,0 monotone_penalty=0.5 lambda_l2=1.0",
    "binary": "objective=binary num_leaves=31 max_bin=63 monotone_constraints=1,-1,0,0,0,0 lambda_l1=0.2 max_delta_step=0.5",
    "with_categorical_and_wide": "objective=regression num_leaves=31 max_bin=600 categorical_feature=4 monotone_constraints=1,-1,0,0,0,0",
    "sampled": "objective=regression num_leaves=15 max_bin=63 monotone_constraints=1,-1,0,0,0,0 feature_fraction=0.7 bagging_freq=1 bagging_fraction=0.7",
}


@pytest.mark.parametrize("case", sorted(CASES))
def test_quantized_models_are_byte_identical_and_monotone(native, oracle, case):
    X, z, rng = mono_data()
    y = (z + rng.normal(0, 0.3, len(z))).astype(np.float32)
    if case == "binary":
        y = (y > 0).astype(np.float32)
    params = CASES[case] + " use_quantized_grad=true stochastic_rounding=false verbosity=-1"
    nd, nb, ob = train_both(oracle, X, y, params, 12)
    ms = nb.model()
    assert strip_parameters(ms) == strip_parameters(ob.save_model_to_string())
    assert np.array_equal(nb.scores(), ob.scores())
    tm = TextModel(ms)
    assert_monotone(tm, np.nan_to_num(X), 0, +1)
    assert_monotone(tm, np.nan_to_num(X), 1, -1)
    used = {int(f) for t in tm.trees if t.num_leaves > 1 for f in t.split_feature}
    assert {0, 1} <= used                                   # the constrained features are split on
    nb.free()
    nd.free()


def test_float_mode_estimator_and_refusals(native, oracle):
    from synapseml_b200.lightgbm.estimators import LightGBMRegressor, LocalDataFrame
    X, z, rng = mono_data()
    y = (z + rng.normal(0, 0.3, len(z))).astype(np.float32)
    params = "objective=regression num_leaves=31 max_bin=255 monotone_constraints=1,-1,0,0,0,0 verbosity=-1"
    nd, nb, ob = train_both(oracle, X, y, params, 8)
    a, b = TextModel(nb.model()), TextModel(ob.save_model_to_string())
    assert a.structure()[:3] == b.structure()[:3]
    assert np.abs(nb.scores() - ob.scores()).max() <= 1e-5 * np.abs(ob.scores()).max()
    nb.free()
    nd.free()
    model = LightGBMRegressor(numIterations=20, numLeaves=31, monotoneConstraints=[1, -1, 0, 0, 0, 0], monotonePenalty=0.5,
                              monotoneConstraintsMethod="basic", samplingMode="fixed").fit(LocalDataFrame.from_arrays(X, y, num_partitions=2))
    text = model.getNativeModel()
    assert "[monotone_constraints: 1,-1,0,0,0,0]" in text and "[monotone_penalty: 0.5]" in text
    tm = TextModel(text)
    assert_monotone(tm, np.nan_to_num(X), 0, +1)
    assert_monotone(tm, np.nan_to_num(X), 1, -1)
    free = LightGBMRegressor(numIterations=20, numLeaves=31, samplingMode="fixed").fit(LocalDataFrame.from_arrays(X, y, num_partitions=2))
    with pytest.raises(AssertionError):
        assert_monotone(TextModel(free.getNativeModel()), np.nan_to_num(X), 0, +1)          # the unconstrained model follows the sine
    for kw in (dict(monotoneConstraintsMethod="advanced"), dict(monotoneConstraints=[1, 0])):
        args = dict(numIterations=2, monotoneConstraints=[1, -1, 0, 0, 0, 0], samplingMode="fixed")
        args.update(kw)
        with pytest.raises(Exception):
            LightGBMRegressor(**args).fit(LocalDataFrame.from_arrays(X, y))
