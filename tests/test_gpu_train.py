"""Whole-path parity through the C ABI: the call sequence of SURVEY.md §3.1-3.2 replayed against the CUDA library and the
oracle.  Quantized mode: byte-identical model text and training scores.  Float mode: identical split features / thresholds
for the first trees and raw scores within 1e-5 relative (the tolerance north_star states)."""
import os

import numpy as np
import pytest

from helpers import NativeBooster, NativeDataset, edge_case_matrix, oracle_dataset, strip_parameters
from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.model_text import TextModel

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
FLOAT_RTOL = 1e-5


def train_both(oracle, X, y, params, iters, weight=None, query=None, micro_batches=(8192,)):
    nd = NativeDataset(X, y, params, weight=weight, query=query, micro_batches=list(micro_batches))
    nb = NativeBooster(nd, params)
    _, dss, _ = oracle_dataset(oracle, X, y, params, weight=weight, query=query)
    ob = oracle.Booster(dss, params)
    fin_n = fin_o = False
    for _ in range(iters):
        fin_n = nb.update()
        fin_o = ob.update_one_iter()
        assert fin_n == fin_o
        if fin_n:
            break
    return nd, nb, ob


QUANT_CASES = {
    "default": "objective=binary num_leaves=31 learning_rate=0.1 max_bin=255 min_data_in_leaf=20",
    "wide_tree": "objective=binary num_leaves=127 learning_rate=0.2 max_bin=63 min_data_in_leaf=5",
    "depth_limited": "objective=binary num_leaves=31 max_depth=3 max_bin=63",
    "regularised": "objective=binary num_leaves=15 lambda_l1=0.5 lambda_l2=2.0 min_gain_to_split=0.1 max_delta_step=0.3 max_bin=63",
    "more_quant_bins": "objective=binary num_leaves=15 num_grad_quant_bins=16 max_bin=63",
    "stochastic": "objective=binary num_leaves=31 max_bin=255 stochastic_rounding=true seed=5",
    "l2": "objective=regression num_leaves=31 max_bin=63 min_data_in_leaf=20",
    "min_hessian": "objective=binary num_leaves=31 max_bin=63 min_sum_hessian_in_leaf=50 min_data_in_leaf=100",
    # posBaggingFraction / negBaggingFraction (LightGBMParams.scala; upstream BalancedBaggingHelper)
    "balanced_bagging": "objective=binary num_leaves=15 max_bin=63 bagging_freq=1 pos_bagging_fraction=0.4 neg_bagging_fraction=0.8 bagging_seed=7",
}


@pytest.mark.parametrize("case", sorted(QUANT_CASES))
def test_quantized_models_are_byte_identical(native, oracle, case):
    params = QUANT_CASES[case] + " use_quantized_grad=true verbosity=-1"
    if "stochastic_rounding" not in params:
        params += " stochastic_rounding=false"
    if case == "l2":
        X, y = S.dense100(40000, f=20)
    else:
        X, y = S.higgs_shape(60000)
    nd, nb, ob = train_both(oracle, X, y, params, 12)
    assert strip_parameters(nb.model()) == strip_parameters(ob.save_model_to_string())
    assert np.array_equal(nb.scores(), ob.scores())
    nb.free()
    nd.free()


def test_quantized_with_weights_nan_features_and_trivial_columns(native, oracle):
    X = edge_case_matrix(20000, 8)
    rng = np.random.Generator(np.random.PCG64(2))
    z = np.nan_to_num(X[:, 0]) + np.isnan(X[:, 3]) * 1.5 - 0.5 * np.nan_to_num(X[:, 2])
    y = (rng.random(len(z)) < 1 / (1 + np.exp(-z))).astype(np.float32)
    w = rng.uniform(0.2, 3.0, len(z)).astype(np.float32)
    params = "objective=binary num_leaves=31 max_bin=63 min_data_in_leaf=20 use_quantized_grad=true stochastic_rounding=false verbosity=-1"
    nd, nb, ob = train_both(oracle, X, y, params, 10, weight=w, micro_batches=(1000, 17))
    ms, mo = nb.model(), ob.save_model_to_string()
    assert strip_parameters(ms) == strip_parameters(mo)
    m = TextModel(ms)
    assert any(((t.decision_type >> 2) & 3 == 2).any() for t in m.trees if t.num_leaves > 1)    # NaN-typed splits were exercised
    assert np.allclose(m.predict_raw(X), nb.scores(), rtol=0, atol=1e-12)                        # default_left routing agrees
    nb.free()
    nd.free()


def test_golden_model_fixture(native):
    X, y = S.higgs_shape(20000)
    params = ("objective=binary num_leaves=15 learning_rate=0.1 max_bin=63 min_data_in_leaf=20 use_quantized_grad=true "
              "stochastic_rounding=false verbosity=-1")
    nd = NativeDataset(X, y, params)
    nb = NativeBooster(nd, params)
    for _ in range(5):
        nb.update()
    assert strip_parameters(nb.model()) == open(os.path.join(GOLDEN, "quant_model.txt")).read()
    nb.free()
    nd.free()


@pytest.mark.parametrize("objective,data", [("binary", "higgs"), ("regression", "dense"), ("binary", "wide")])
def test_float_mode_first_trees_and_scores(native, oracle, objective, data):
    if data == "higgs":
        X, y = S.higgs_shape(100000)
    elif data == "dense":
        X, y = S.dense100(50000, f=40)
    else:
        X, y = S.epsilon_shape(8000, f=200)
    params = "objective=%s num_leaves=31 learning_rate=0.1 max_bin=255 min_data_in_leaf=20 verbosity=-1" % objective
    nd, nb, ob = train_both(oracle, X, y, params, 10)
    mn, mo = TextModel(nb.model()), TextModel(ob.save_model_to_string())
    sn, so = mn.structure(), mo.structure()
    assert sn[:3] == so[:3]                       # identical split features, thresholds and topology for the first trees
    same = sum(a == b for a, b in zip(sn, so))
    assert same >= 8, same
    a, b = nb.scores(), ob.scores()
    assert np.abs(a - b).max() <= FLOAT_RTOL * np.abs(b).max()
    nb.free()
    nd.free()


def test_lambdarank_float_and_quantized(native, oracle):
    X, rel, qid = S.mslr_shape(30000, f=40)
    base = "objective=lambdarank num_leaves=31 max_bin=63 min_data_in_leaf=20 max_position=20 eval_at=1,2,3,4,5 verbosity=-1"
    nd, nb, ob = train_both(oracle, X, rel, base, 1, query=qid)
    g, h = nb.gradients()
    og, oh = ob.gradients()
    assert np.array_equal(g, og) and np.array_equal(h, oh)      # scores are identical before the first tree: bit-exact lambdas
    for _ in range(5):
        nb.update()
        ob.update_one_iter()
    a, b = nb.scores(), ob.scores()
    assert np.abs(a - b).max() <= 1e-4 * max(np.abs(b).max(), 1e-12)
    assert TextModel(nb.model()).structure()[:2] == TextModel(ob.save_model_to_string()).structure()[:2]
    nb.free()
    nd.free()
    # quantized lambdarank: first-iteration gradients are bit-exact (per-document float32 accumulation replays the CPU order)
    q = base + " use_quantized_grad=true stochastic_rounding=false"
    nd, nb, ob = train_both(oracle, X, rel, q, 1, query=qid)
    g, h = nb.gradients()
    og, oh = ob.gradients()
    assert np.array_equal(g, og) and np.array_equal(h, oh)
    assert strip_parameters(nb.model()) == strip_parameters(ob.save_model_to_string())
    nb.free()
    nd.free()


def test_training_stops_when_no_leaf_can_split(native, oracle):
    X = np.array([[1.0]] * 30 + [[2.0]] * 30)
    y = np.array([0.0] * 30 + [1.0] * 30, dtype=np.float32)
    params = "objective=regression learning_rate=1.0 min_data_in_leaf=1 min_data_in_bin=1 num_leaves=4 use_quantized_grad=false verbosity=-1"
    nd, nb, ob = train_both(oracle, X, y, params, 5)
    mn, mo = TextModel(nb.model()), TextModel(ob.save_model_to_string())
    assert len(mn.trees) == len(mo.trees) == 1            # the second iteration finds no split, is popped and reports finished
    assert mn.trees[0].leaf_value.tolist() == [0.0, 1.0] == mo.trees[0].leaf_value.tolist()
    assert np.array_equal(nb.scores(), y.astype(np.float64))
    nb.free()
    nd.free()


@pytest.mark.parametrize("boost_from_average", ["true", "false"])
def test_single_class_shard_gives_constant_model(native, oracle, boost_from_average):
    """GBDT::TrainOneIter's "only add default score one-time" branch: no tree can be trained, the model is one constant tree holding
    the automatic initial score, and that score enters the training scores exactly once (with or without boost_from_average)."""
    X, _ = S.higgs_shape(2000)
    y = np.ones(2000, dtype=np.float32)
    params = "objective=binary num_leaves=7 boost_from_average=%s verbosity=-1" % boost_from_average
    nd, nb, ob = train_both(oracle, X, y, params, 2)
    assert strip_parameters(nb.model()) == strip_parameters(ob.save_model_to_string())
    m = TextModel(nb.model())
    assert len(m.trees) == 1 and m.trees[0].num_leaves == 1
    expected = np.log((1 - 1e-15) / 1e-15)        # pavg clamped at 1 - kEpsilon (float32 1e-15)
    assert np.allclose(m.trees[0].leaf_value, expected, rtol=1e-3)
    assert np.array_equal(nb.scores(), ob.scores())
    assert np.allclose(nb.scores(), m.trees[0].leaf_value[0], rtol=0, atol=0)      # once, not twice
    nb.free()
    nd.free()


def test_constant_features_give_constant_tree_of_the_label_mean(native, oracle):
    """No usable feature at all (every column trivial): regression starts from the label mean and stops; scores hold it once."""
    X = np.full((500, 3), 2.5)
    y = np.arange(500, dtype=np.float32)
    params = "objective=regression num_leaves=7 min_data_in_leaf=1 verbosity=-1"
    nd, nb, ob = train_both(oracle, X, y, params, 2)
    assert strip_parameters(nb.model()) == strip_parameters(ob.save_model_to_string())
    assert np.array_equal(nb.scores(), ob.scores())
    assert np.allclose(nb.scores(), y.astype(np.float64).mean())
    nb.free()
    nd.free()


def test_full_size_properties(native):
    """BASELINE-sized run (scaled to what a test can afford: 4M x 28, quantized): size-independent properties."""
    X, y = S.higgs_shape(4_000_000)
    params = "objective=binary num_leaves=31 max_bin=255 min_data_in_leaf=20 use_quantized_grad=true verbosity=-1"
    nd = NativeDataset(X, y, params, micro_batches=[1 << 16])
    nb = NativeBooster(nd, params)
    for _ in range(5):
        assert not nb.update()
    hist = nb.root_histogram()
    qg, qh, _ = nb.quant_gradients()
    assert (hist[:, :, 0].sum(axis=1) == qg.astype(np.int64).sum()).all()      # every feature's bins sum to the leaf totals
    assert (hist[:, :, 1].sum(axis=1) == qh.astype(np.int64).sum()).all()
    m = TextModel(nb.model())
    for t in m.trees:
        assert t.leaf_count.sum() == len(y) and t.num_leaves == 31              # the partition loses no row
        assert (t.internal_count[0] == len(y))
    sub = np.random.Generator(np.random.PCG64(0)).choice(len(y), 200000, replace=False)
    assert np.allclose(m.predict_raw(X[sub]), nb.scores()[sub], rtol=0, atol=1e-12)   # score update == tree traversal
    nb.free()
    nd.free()


@pytest.mark.parametrize("mode", ["quantized", "float"])
def test_both_split_search_forms_give_the_same_model(native, monkeypatch, mode):
    """The split search runs as a 256-thread CTA per (feature, child, leaf) on narrow data and as a warp per unit on wide data
    (LGB_B200_FIND_WARP forces either): same thresholds, gains and tie-breaks, so the same model, on narrow and on wide data."""
    params = "objective=binary num_leaves=31 max_bin=255 min_data_in_leaf=20 verbosity=-1"
    if mode == "quantized":
        params += " use_quantized_grad=true stochastic_rounding=false"
    Xn, yn = S.higgs_shape(60000)
    Xw, yw = S.epsilon_shape(6000, f=300)
    Xw[::7, 3] = np.nan                      # a NaN bin: the forward scan runs too
    for X, y in ((Xn, yn), (Xw, yw)):
        models = []
        for flag in ("0", "1"):
            monkeypatch.setenv("LGB_B200_FIND_WARP", flag)
            nd = NativeDataset(X, y, params, micro_batches=[8192])
            nb = NativeBooster(nd, params)
            for _ in range(6):
                nb.update()
            models.append(strip_parameters(nb.model()))
            nb.free()
            nd.free()
        assert models[0] == models[1]


@pytest.mark.parametrize("mode", ["quantized", "float"])
def test_zero_as_missing(native, oracle, mode, monkeypatch):
    """zeroAsMissing (LightGBMParams.scala:176): zeros and NaNs are the missing value; the split search leaves the zero bin out of
    both scans (SKIP_DEFAULT_BIN), rows of that bin follow default_left in the partition, the score pass and the predictor."""
    rng = np.random.Generator(np.random.PCG64(11))
    n = 40000
    X = rng.standard_normal((n, 8))
    X[:, 0] = np.where(rng.random(n) < 0.5, 0.0, rng.exponential(1.0, n))                 # zero is the most frequent bin and bin 0
    X[:, 1] = np.where(rng.random(n) < 0.3, 0.0, rng.standard_normal(n))                  # zero in the middle of the range
    X[:, 2] = np.where(rng.random(n) < 0.2, np.nan, np.where(rng.random(n) < 0.3, 0.0, rng.standard_normal(n) - 2.0))   # NaN counts as zero
    X[:, 3] = np.where(rng.random(n) < 0.85, -1.5, rng.standard_normal(n))                # most frequent bin 0 is NOT the zero bin
    X[:, 4] = rng.integers(0, 3, n).astype(np.float64)
    z = 1.2 * (X[:, 0] == 0) - 0.8 * (np.nan_to_num(X[:, 1]) == 0) + 0.7 * np.nan_to_num(X[:, 2]) + 0.5 * X[:, 3] + 0.4 * X[:, 5]
    y = (rng.random(n) < 1 / (1 + np.exp(-z))).astype(np.float32)
    params = "objective=binary num_leaves=31 max_bin=63 min_data_in_leaf=20 zero_as_missing=true verbosity=-1"
    if mode == "quantized":
        params += " use_quantized_grad=true stochastic_rounding=false"
    nd, nb, ob = train_both(oracle, X, y, params, 10, micro_batches=(4096,))
    ms, mo = nb.model(), ob.save_model_to_string()
    m = TextModel(ms)
    if mode == "quantized":
        assert strip_parameters(ms) == strip_parameters(mo)
        assert np.array_equal(nb.scores(), ob.scores())
    else:
        assert m.structure()[:4] == TextModel(mo).structure()[:4]
        assert np.abs(nb.scores() - ob.scores()).max() <= FLOAT_RTOL * np.abs(ob.scores()).max()
    assert any(((t.decision_type >> 2) & 3 == 1).any() for t in m.trees if t.num_leaves > 1)      # Zero-typed splits were exercised
    assert np.allclose(m.predict_raw(X), nb.scores(), rtol=0, atol=1e-12)                          # bins and raw values route alike
    nb.free()
    # the warp form of the split search skips the zero bin in the same way
    monkeypatch.setenv("LGB_B200_FIND_WARP", "1")
    nw = NativeBooster(nd, params)
    for _ in range(10):
        nw.update()
    assert strip_parameters(nw.model()) == strip_parameters(ms)
    nw.free()
    nd.free()


def test_reset_parameter_changes_the_learning_rate_of_later_trees(native):
    """LGBM_BoosterResetParameter("learning_rate=...") is how the reference's delegate changes the learning rate between iterations
    (TrainUtils.scala:98-135 -> LightGBMBooster.resetParameter): the trees that follow are shrunk by the new rate."""
    import ctypes as C
    from synapseml_b200.lightgbm.native import validate
    X, y = S.higgs_shape(30000)
    params = "objective=binary num_leaves=15 max_bin=63 learning_rate=0.1 use_quantized_grad=true stochastic_rounding=false verbosity=-1"
    nd = NativeDataset(X, y, params)
    nb = NativeBooster(nd, params)
    for _ in range(3):
        nb.update()
    validate(nb.lib.LGBM_BoosterResetParameter(nb.handle, b"learning_rate=0.5"))
    for _ in range(2):
        nb.update()
    text = nb.model()
    shr = [float(l.split("=")[1]) for l in text.split("\n") if l.startswith("shrinkage=")]
    assert shr == [1.0, 0.1, 0.1, 0.5, 0.5]               # (the first tree carries the initial score: AddBias resets it to 1)
    m = TextModel(text)
    assert np.allclose(m.predict_raw(X), nb.scores(), rtol=0, atol=1e-12)
    assert np.abs(m.trees[3].leaf_value).max() > 2 * np.abs(m.trees[2].leaf_value).max()
    # a model trained at 0.5 throughout differs, one trained at 0.1 throughout agrees on the first three trees
    nb2 = NativeBooster(nd, params)
    for _ in range(3):
        nb2.update()
    m2 = TextModel(nb2.model())
    assert m2.structure() == m.structure()[:3] and all(np.array_equal(a.leaf_value, b.leaf_value) for a, b in zip(m2.trees, m.trees))
    nb2.free(); nb.free(); nd.free()
