"""Column and row sampling (feature_fraction, bagging_fraction / bagging_freq) through the C ABI against
the oracle: the same LCG streams pick the same features per tree and the same bag per re-bagging, so quantized models stay
byte-identical and float models keep their structure."""
import numpy as np
import pytest

from helpers import NativeBooster, NativeDataset, edge_case_matrix, oracle_dataset, strip_parameters
from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.model_text import TextModel
from test_gpu_train import FLOAT_RTOL, train_both

pytestmark = pytest.mark.gpu

QUANT = " use_quantized_grad=true stochastic_rounding=false verbosity=-1"
CASES = {
    "feature_fraction": "objective=binary num_leaves=31 max_bin=63 feature_fraction=0.6",
    "feature_fraction_seeded": "objective=binary num_leaves=15 max_bin=63 feature_fraction=0.3 feature_fraction_seed=11",
    "feature_fraction_from_seed": "objective=binary num_leaves=15 max_bin=63 feature_fraction=0.5 seed=42",
    "bagging": "objective=binary num_leaves=31 max_bin=63 bagging_fraction=0.7 bagging_freq=1",
    "bagging_every_third": "objective=binary num_leaves=31 max_bin=255 bagging_fraction=0.4 bagging_freq=3 bagging_seed=9",
    "both_l2": "objective=regression num_leaves=31 max_bin=63 bagging_fraction=0.8 bagging_freq=2 feature_fraction=0.7",
    # more than 32 features: the split search's own last-unit reduction instead of the replay kernel's
    "both_l2_wide": "objective=regression num_leaves=31 max_bin=63 bagging_fraction=0.8 bagging_freq=1 feature_fraction=0.5",
    # featureFractionByNode (LightGBMParams.scala; ColSampler::GetByNode): a sample per searched leaf, from the tree's own sample
    # when there is one; trees that fill up, trees that stop early (the number of samples a tree uses decides where the next
    # tree's stream starts), children that are never searched (max_depth)
    "bynode": "objective=binary num_leaves=31 max_bin=63 feature_fraction_bynode=0.5",
    "bynode_and_bytree": "objective=binary num_leaves=15 max_bin=63 feature_fraction=0.7 feature_fraction_bynode=0.6 feature_fraction_seed=5",
    "bynode_early_stop": "objective=binary num_leaves=63 max_bin=63 feature_fraction_bynode=0.4 min_gain_to_split=40 min_data_in_leaf=200",
    "bynode_max_depth": "objective=binary num_leaves=31 max_depth=3 max_bin=63 feature_fraction_bynode=0.5 bagging_fraction=0.8 bagging_freq=1",
    "bynode_l2_wide": "objective=regression num_leaves=31 max_bin=63 feature_fraction_bynode=0.3",
}


@pytest.mark.parametrize("case", sorted(CASES))
def test_quantized_models_with_sampling_are_byte_identical(native, oracle, case):
    params = CASES[case] + QUANT
    if case.startswith("both_l2") or case == "bynode_l2_wide":
        X, y = S.dense100(40000, f=48 if case.endswith("wide") else 20)
    else:
        X, y = S.higgs_shape(50000 + 317)         # a ragged last 1024-row bagging block
    nd, nb, ob = train_both(oracle, X, y, params, 10)
    ms, mo = nb.model(), ob.save_model_to_string()
    assert strip_parameters(ms) == strip_parameters(mo)
    assert np.array_equal(nb.scores(), ob.scores())          # rows outside the bag are scored through the tree as well
    if case == "bynode_early_stop":
        sizes = [t.num_leaves for t in TextModel(ms).trees]
        assert min(sizes) < 63 and len(set(sizes)) > 1                   # trees that stopped before they were full
    if "feature_fraction" in case:
        used = [set(t.split_feature[:t.num_leaves - 1].tolist()) for t in TextModel(ms).trees]
        assert len(set.union(*used)) > max(len(u) for u in used)      # the trees drew different column samples
    if "bagging" in case:
        t0 = TextModel(ms).trees[0]
        assert t0.internal_count[0] < len(X)                           # the root counts the bag, not the data
    nb.free()
    nd.free()


def test_float_mode_with_sampling_keeps_structure(native, oracle):
    X, y = S.higgs_shape(80000)
    params = "objective=binary num_leaves=31 max_bin=255 bagging_fraction=0.6 bagging_freq=1 feature_fraction=0.8 verbosity=-1"
    nd, nb, ob = train_both(oracle, X, y, params, 8)
    sn, so = TextModel(nb.model()).structure(), TextModel(ob.save_model_to_string()).structure()
    assert sn[:3] == so[:3]
    assert sum(a == b for a, b in zip(sn, so)) >= 6
    a, b = nb.scores(), ob.scores()
    assert np.abs(a - b).max() <= FLOAT_RTOL * np.abs(b).max()
    nb.free()
    nd.free()


def test_bagging_with_nan_features_and_weights(native, oracle):
    X = edge_case_matrix(20000, 8)
    rng = np.random.Generator(np.random.PCG64(2))
    z = np.nan_to_num(X[:, 0]) + np.isnan(X[:, 3]) * 1.5 - 0.5 * np.nan_to_num(X[:, 2])
    y = (rng.random(len(z)) < 1 / (1 + np.exp(-z))).astype(np.float32)
    w = rng.uniform(0.2, 3.0, len(z)).astype(np.float32)
    params = "objective=binary num_leaves=31 max_bin=63 bagging_fraction=0.5 bagging_freq=1 feature_fraction=0.75" + QUANT
    nd, nb, ob = train_both(oracle, X, y, params, 8, weight=w, micro_batches=(1000, 17))
    assert strip_parameters(nb.model()) == strip_parameters(ob.save_model_to_string())
    assert np.array_equal(nb.scores(), ob.scores())
    nb.free()
    nd.free()


def test_unimplemented_sampling_modes_are_refused(native):
    X, y = S.higgs_shape(5000)
    base = "objective=binary num_leaves=7 max_bin=63 verbosity=-1 "
    nd = NativeDataset(X, y, base)
    for extra in ("feature_fraction_bynode=1.5", "bagging_freq=1 pos_bagging_fraction=1.5", "monotone_constraints=1,0,-1 monotone_constraints_method=advanced",
                  "extra_trees=true", "feature_fraction=0"):
        with pytest.raises(Exception):
            NativeBooster(nd, base + extra)
    nd.free()
