"""Custom objective (TrainUtils.scala:80-86, LightGBMBooster.scala:331-392): LGBM_BoosterGetPredict hands
the raw training scores out, LGBM_BoosterUpdateOneIterCustom takes the caller's gradients.  Fed with the gradients the built-in
objective computed, the custom path must grow the very same trees."""
import ctypes as C

import numpy as np
import pytest

from helpers import NativeBooster, NativeDataset
from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LocalDataFrame
from synapseml_b200.lightgbm.model_text import TextModel
from synapseml_b200.lightgbm.native import validate

pytestmark = pytest.mark.gpu


def update_custom(nb, g, h):
    fin = C.c_int()
    validate(nb.lib.LGBM_BoosterUpdateOneIterCustom(nb.handle, g.ctypes.data_as(C.POINTER(C.c_float)), h.ctypes.data_as(C.POINTER(C.c_float)),
                                                    C.byref(fin)))
    return fin.value == 1


def tree_blocks(model_str):
    return model_str.split("end of trees")[0].split("Tree=0\n", 1)[1]


@pytest.mark.parametrize("mode", ["quantized", "float"])
def test_builtin_gradients_through_the_custom_entry_point_grow_the_same_trees(native, mode):
    X, y = S.higgs_shape(50000)
    common = " num_leaves=31 max_bin=63 min_data_in_leaf=20 boost_from_average=false verbosity=-1"
    if mode == "quantized":
        common += " use_quantized_grad=true stochastic_rounding=false"
    nd = NativeDataset(X, y, "objective=binary" + common)
    a = NativeBooster(nd, "objective=binary" + common)
    b = NativeBooster(nd, "objective=custom" + common)
    for _ in range(8):
        assert not a.update()
        g, h = a.gradients()                    # the gradients this iteration of `a` was grown from
        assert not update_custom(b, g, h)
    ma, mb = a.model(), b.model()
    assert tree_blocks(ma) == tree_blocks(mb)
    assert np.array_equal(a.scores(), b.scores())
    assert "objective=binary" in ma and "\nobjective=" not in mb.split("Tree=0")[0]      # no objective function, no objective line
    with pytest.raises(Exception):               # a custom-objective booster has no gradients of its own
        b.update()
    a.free()
    b.free()
    nd.free()


def test_user_gradients_give_the_oracles_model(native, oracle):
    """Gradients that no built-in objective produces (the reference test's log-likelihood, computed in numpy from the raw scores
    LGBM_BoosterGetPredict returns) go to the CUDA library and to the oracle alike: quantized models are byte-identical."""
    from helpers import oracle_dataset, strip_parameters
    X, y = S.higgs_shape(40000)
    params = ("objective=custom num_leaves=31 max_bin=63 min_data_in_leaf=20 use_quantized_grad=true stochastic_rounding=false "
              "bagging_fraction=0.8 bagging_freq=1 verbosity=-1")
    nd = NativeDataset(X, y, params)
    nb = NativeBooster(nd, params)
    _, dss, _ = oracle_dataset(oracle, X, y, params)
    ob = oracle.Booster(dss, params)
    for _ in range(8):
        s = nb.scores()
        assert np.array_equal(s, ob.scores())
        prob = 1.0 / (1.0 + np.exp(-s))
        g, h = (prob - y).astype(np.float32), (prob * (1.0 - prob)).astype(np.float32)
        assert update_custom(nb, g, h) == ob.update_one_iter_custom(g, h)
    assert strip_parameters(nb.model()) == strip_parameters(ob.save_model_to_string())
    nb.free()
    nd.free()


class LogLikelihood:
    """The reference test's FObjTrait (VerifyLightGBMClassifierStream.scala:57-72)."""

    def get_gradient(self, predictions, training_data):
        labels = training_data.get_label()
        prob = 1.0 / (1.0 + np.exp(-predictions[:, 0]))
        return (prob - labels).astype(np.float32), (prob * (1.0 - prob)).astype(np.float32)


def test_estimator_with_fobj_learns_and_scores_raw(native):
    X, y = S.higgs_shape(40000)
    df = LocalDataFrame.from_arrays(X, y, num_partitions=1)
    model = LightGBMClassifier(numIterations=30, numLeaves=15, fobj=LogLikelihood()).fit(df)
    ms = model.getNativeModel()
    assert "[objective: custom]" in ms and len(TextModel(ms).trees) == 30
    raw = TextModel(ms).predict_raw(X)
    order = np.argsort(raw, kind="stable")
    r = np.empty(len(raw)); r[order] = np.arange(1, len(raw) + 1)
    pos = y > 0
    auc = (r[pos].sum() - pos.sum() * (pos.sum() + 1) / 2) / (pos.sum() * (~pos).sum())
    assert auc > 0.75
    builtin = LightGBMClassifier(numIterations=30, numLeaves=15).fit(df)
    # same loss, same data: the custom model tracks the built-in one (it starts from 0 instead of the label average)
    rb = TextModel(builtin.getNativeModel()).predict_raw(X)
    assert np.corrcoef(raw, rb)[0, 1] > 0.98


def test_custom_objective_with_several_classes(native, oracle):
    """FObjTrait with a multiclass label (TrainUtils.scala:80-86 hands the [rows][classes] scores to the user and the returned arrays
    — one value per row and class, class after class — to LGBM_BoosterUpdateOneIterCustom): num_class trees per iteration, no
    objective line in the model.  Fed with the built-in softmax gradients the custom path grows the very same trees; with gradients
    of its own it matches the oracle byte for byte."""
    from helpers import oracle_dataset, strip_parameters
    X, _ = S.higgs_shape(40000)
    rng = np.random.Generator(np.random.PCG64(0))
    y = np.digitize(X[:, 0] + 0.5 * X[:, 2] + 0.3 * rng.standard_normal(len(X)), [-0.6, 0.7]).astype(np.float32)
    common = " num_class=3 num_leaves=15 max_bin=63 min_data_in_leaf=20 boost_from_average=false use_quantized_grad=true stochastic_rounding=false verbosity=-1"
    nd = NativeDataset(X, y, "objective=multiclass" + common)
    a = NativeBooster(nd, "objective=multiclass" + common)
    b = NativeBooster(nd, "objective=custom" + common)
    assert b.num_classes() == 3
    for _ in range(5):
        assert not a.update()
        g, h = a.gradients()                    # [3][n], class-major
        assert g.size == 3 * len(y)
        assert not update_custom(b, g, h)
    ma, mb = a.model(), b.model()
    assert tree_blocks(ma) == tree_blocks(mb) and np.array_equal(a.scores(), b.scores())
    assert "num_class=3" in mb and "num_tree_per_iteration=3" in mb and "\nobjective=" not in mb.split("Tree=0")[0]
    a.free(); b.free()
    # user gradients (one-vs-rest log-likelihood in numpy) against the oracle
    params = "objective=custom" + common
    nb = NativeBooster(nd, params)
    _, dss, _ = oracle_dataset(oracle, X, y, params)
    ob = oracle.Booster(dss, params)
    for _ in range(4):
        s = nb.scores().reshape(3, -1)
        assert np.array_equal(s.reshape(-1), ob.scores())
        prob = 1.0 / (1.0 + np.exp(-s))
        onehot = (np.arange(3)[:, None] == y[None, :].astype(int))
        g, h = (prob - onehot).astype(np.float32).reshape(-1), (prob * (1.0 - prob)).astype(np.float32).reshape(-1)
        assert update_custom(nb, g, h) == ob.update_one_iter_custom(g, h)
    assert strip_parameters(nb.model()) == strip_parameters(ob.save_model_to_string())
    nb.free()
    nd.free()

    # through the estimator: a multiclass FObjTrait
    class OneVsRest:
        def get_gradient(self, predictions, training_data):
            labels = training_data.get_label().astype(int)
            assert predictions.shape == (len(labels), 3)
            prob = 1.0 / (1.0 + np.exp(-predictions))
            hot = np.arange(3)[None, :] == labels[:, None]
            return (prob - hot).T.astype(np.float32).reshape(-1), (prob * (1.0 - prob)).T.astype(np.float32).reshape(-1)

    model = LightGBMClassifier(objective="multiclass", fobj=OneVsRest(), numIterations=10, numLeaves=15, samplingMode="fixed").fit(
        LocalDataFrame.from_arrays(X, y, num_partitions=2))
    out = model.transform(X[:5000])
    assert out["rawPrediction"].shape == (5000, 3) and (out["prediction"] == y[:5000]).mean() > 0.8
