"""Host-side mirror of the reference's Scala layer: the pure-logic known answers the reference itself tests."""
import numpy as np
import pytest

from synapseml_b200.lightgbm.dataset import GroupIdManager, SampledData, count_cardinality
from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LightGBMRanker, LightGBMRegressor, LocalDataFrame
from synapseml_b200.lightgbm.params import (ClassifierTrainParams, GeneralParams, ObjectiveParams, ParamsStringBuilder,
                                            QuantizedGradParams, RankerTrainParams, dataset_creation_params)


def test_count_cardinality_reference_vector():
    # lightgbm/src/test/scala/.../split2/VerifyLightGBMRankerStream.scala:76-89
    assert count_cardinality([1, 1, 2, 2, 2, 3]) == [2, 3, 1]
    assert count_cardinality([1 << 40, 1 << 40, 2, 2, 2, 3]) == [2, 3, 1]
    assert count_cardinality(["a", "a", "b", "b", "b", "c"]) == [2, 3, 1]


def test_params_string_builder_append_if_not_there():
    # core/src/test/scala/.../core/utils/VerifyParamsStringBuilder.scala
    sb = ParamsStringBuilder(prefix="--", delimiter=" ")
    sb.append("--first_param 1").append_param_value_if_not_there("first_param", 2).append_param_value_if_not_there("second", 3.5)
    sb.append_param_flag_if_not_there("flag").append_param_flag_if_not_there("flag")
    sb.append_param_list_if_not_there("lst", [1, 2, 3]).append_param_value_if_not_there("none", None)
    assert sb.result() == "--first_param 1 --second 3.5 --flag --lst 1,2,3"


def test_booster_param_string_first_writer_wins():
    p = ClassifierTrainParams(passThroughArgs="max_bin=63 learning_rate=0.3",
                              generalParams=GeneralParams(numLeaves=31, maxBin=255, learningRate=0.1, minSumHessianInLeaf=1e-4, numIterations=7),
                              objectiveParams=ObjectiveParams("binary"), quantizedParams=QuantizedGradParams(True))
    s = p.to_string()
    assert s.startswith("max_bin=63 learning_rate=0.3 is_pre_partition=True boosting_type=gbdt tree_learner=data_parallel num_leaves=31")
    assert s.count("max_bin=") == 1 and s.count("learning_rate=") == 1
    assert "min_sum_hessian_in_leaf=1.0E-4" in s          # JVM Double.toString
    assert "num_iterations=7" in s and "objective=binary" in s and "use_quantized_grad=true" in s
    assert s.endswith("is_unbalance=false boost_from_average=true")
    r = RankerTrainParams(objectiveParams=ObjectiveParams("lambdarank")).to_string()
    assert r.endswith("eval_at=1,2,3,4,5 max_position=20")


def test_dataset_creation_params():
    assert dataset_creation_params(255, 200000, 20, 7) == "is_pre_partition=True max_bin=255 bin_construct_sample_cnt=200000 min_data_in_leaf=20 num_threads=7"
    assert dataset_creation_params(255, 200000, 20, 7, data_random_seed=3).endswith("data_random_seed=3")


def test_sampled_data_layout():
    # SampledData.scala:43-77: column-major, zeros skipped, NaN kept, row index recorded
    rows = np.array([[0.0, 1.5, np.nan], [2.0, 0.0, 0.0], [3.0, 0.0, 4.0]])
    sd = SampledData(3, 3)
    for i, r in enumerate(rows):
        sd.push_row(r, i)
    assert sd.row_counts.tolist() == [2, 1, 2]
    assert sd.sample_data[0][:2].tolist() == [2.0, 3.0] and sd.sample_indexes[0][:2].tolist() == [1, 2]
    assert sd.sample_data[1][:1].tolist() == [1.5] and sd.sample_indexes[1][:1].tolist() == [0]
    assert np.isnan(sd.sample_data[2][0]) and sd.sample_data[2][1] == 4.0 and sd.sample_indexes[2][:2].tolist() == [0, 2]
    sd2 = SampledData(3, 3)
    sd2.push_rows(rows)
    assert sd2.row_counts.tolist() == sd.row_counts.tolist()
    for j in range(3):
        k = sd.row_counts[j]
        assert np.array_equal(sd.sample_data[j][:k], sd2.sample_data[j][:k], equal_nan=True)
        assert np.array_equal(sd.sample_indexes[j][:k], sd2.sample_indexes[j][:k])


def test_sampled_data_blocks_equal_row_by_row_pushes():
    # push_rows over several blocks and calls: columns without zeros, with zeros, all zero, with NaN and -0.0
    rng = np.random.Generator(np.random.PCG64(4))
    rows = rng.normal(size=(700, 6))
    rows[:, 1] = np.where(rng.random(700) < 0.5, 0.0, rows[:, 1])
    rows[:, 2] = 0.0
    rows[rng.random(700) < 0.1, 3] = np.nan
    rows[rng.random(700) < 0.3, 4] = -0.0
    rows[300:, 5] = np.where(rng.random(400) < 0.2, 0.0, rows[300:, 5])      # full in the first blocks only
    one = SampledData(700, 6)
    for i, r in enumerate(rows):
        one.push_row(r, i)
    blocks = SampledData(700, 6)
    blocks.push_rows(rows[:450], block=128)
    blocks.push_rows(rows[450:], block=100)
    assert one.row_counts.tolist() == blocks.row_counts.tolist() and one.row_counts[2] == 0 and one.row_counts[0] == 700
    for j in range(6):
        k = one.row_counts[j]
        assert np.array_equal(one.sample_data[j][:k], blocks.sample_data[j][:k], equal_nan=True)
        assert np.array_equal(one.sample_indexes[j][:k], blocks.sample_indexes[j][:k])


def test_group_id_manager():
    g = GroupIdManager()
    assert [g.get_unique_index_for_id(v) for v in ["q7", "q7", "q2", "q7", 5]] == [0, 0, 1, 0, 2]


def test_estimator_params_and_defaults():
    c = LightGBMClassifier().setNumLeaves(15).setExecutionMode("bulk").setUseQuantizedGrad(True)
    assert c.getNumLeaves() == 15 and c.getDataTransferMode() == "bulk" and c.getMaxBin() == 255 and c.getObjective() == "binary"
    assert c.getNumIterations() == 100 and c.getLearningRate() == 0.1 and c.getMinDataInLeaf() == 20
    s = c.get_train_params(1, 4).to_string()
    assert "num_leaves=15" in s and "max_bin" not in s and "num_threads=3" in s and "use_quantized_grad=true" in s
    assert LightGBMRegressor().getObjective() == "regression" and LightGBMRanker().getObjective() == "lambdarank"
    assert LightGBMRanker().getMaxPosition() == 20
    with pytest.raises(ValueError):
        LightGBMClassifier().setDataTransferMode("neither")
    with pytest.raises(AttributeError):
        LightGBMClassifier().setNoSuchParam(1)


def test_sampling_modes():
    X = np.arange(50000 * 2, dtype=np.float64).reshape(50000, 2)
    df = LocalDataFrame.from_arrays(X, np.zeros(50000), num_partitions=3)
    est = LightGBMClassifier(binSampleCount=1000, samplingMode="fixed")
    assert np.array_equal(est.get_sampled_rows(df, 50000), X[:1000])
    est = LightGBMClassifier(binSampleCount=1000, samplingMode="global")
    s = est.get_sampled_rows(df, 50000)
    assert len(s) == 1000
    # "subset": limit(min(samplingSubsetSize, numSamples)).sample(fraction) with no final limit -> ~fraction * 1000 rows
    est = LightGBMClassifier(binSampleCount=1000, samplingMode="subset")
    s = est.get_sampled_rows(df, 50000)
    assert 150 < len(s) < 300 and s[:, 0].max() < 2000


def test_ranker_sorts_within_partitions_and_rejects_bad_group_column():
    df = LocalDataFrame([{"features": np.arange(8.0).reshape(4, 2), "label": np.arange(4.0), "group": np.array([3, 1, 3, 1])}])
    out = LightGBMRanker().preprocess_data(df).partitions[0]
    assert out["group"].tolist() == [1, 1, 3, 3] and out["label"].tolist() == [1.0, 3.0, 0.0, 2.0]
    bad = LocalDataFrame([{"features": np.zeros((2, 2)), "label": np.zeros(2), "group": np.array([0.5, 1.5])}])
    with pytest.raises(ValueError):
        LightGBMRanker().preprocess_data(bad)


def test_round_two_params_reach_the_native_string():
    """CategoricalParams, DartModeParams, SeedParams, monotone constraints and the categorical slot resolution
    (LightGBMBase.scala:167-198, params/TrainParams.scala)."""
    c = LightGBMClassifier(boostingType="dart", dropRate=0.2, maxDrop=7, skipDrop=0.3, uniformDrop=True, xGBoostDartMode=True, dropSeed=11,
                           featureFractionSeed=5, extraSeed=6, baggingSeed=9, categoricalSlotIndexes=[4], categoricalSlotNames=["b", "a"],
                           slotNames=["a", "b", "c", "d", "e"], minDataPerGroup=50, maxCatThreshold=8, catl2=2.0, catSmooth=5.0, maxCatToOnehot=6,
                           monotoneConstraints=[1, 0, -1, 0, 0], monotoneConstraintsMethod="basic", monotonePenalty=0.5, maxBinByFeature=[3, 4, 5, 6, 7])
    assert c.get_categorical_indexes() == [4, 1, 0]
    s = c.get_train_params(1, 2).to_string()
    for frag in ("boosting_type=dart", "drop_rate=0.2", "max_drop=7", "skip_drop=0.3", "uniform_drop=true", "xgboost_dart_mode=true", "drop_seed=11",
                 "feature_fraction_seed=5", "extra_seed=6", "bagging_seed=9", "categorical_feature=4,1,0", "min_data_per_group=50",
                 "max_cat_threshold=8", "cat_l2=2.0", "cat_smooth=5.0", "max_cat_to_onehot=6", "monotone_constraints=1,0,-1,0,0",
                 "monotone_constraints_method=basic", "monotone_penalty=0.5", "max_bin_by_feature=3,4,5,6,7"):
        assert frag in s.split(), (frag, s)
    plain = LightGBMClassifier(dropRate=0.2, dropSeed=11, featureFractionByNode=0.5).get_train_params(1, 1).to_string()
    assert "drop_rate" not in plain and "drop_seed" not in plain              # DART parameters only travel with boosting_type=dart
    assert "feature_fraction_bynode" not in plain                             # declared by the reference, never written (BaseTrainParams.scala)
    r = LightGBMRanker(objective="rank_xendcg", objectiveSeed=3).get_train_params(1, 1).to_string()
    assert "objective=rank_xendcg" in r and "objective_seed=3" in r
    assert "objective_seed" not in LightGBMRanker(objectiveSeed=3).get_train_params(1, 1).to_string()
    with pytest.raises(ValueError):
        LightGBMClassifier(categoricalSlotNames=["a"]).get_categorical_indexes()      # names need slotNames


def test_column_name_params_pick_the_columns():
    X, y, w = np.zeros((6, 2)), np.arange(6.0), np.ones(6)
    df = LocalDataFrame([{"x": X, "target": y, "wt": w, "base": y * 0.1}])
    est = LightGBMRegressor(labelCol="target", featuresCol="x", weightCol="wt", initScoreCol="base")
    p = est._named_columns(df).partitions[0]
    assert p["label"] is y and p["features"] is X and p["weight"] is w and np.array_equal(p["initScore"], y * 0.1)
    with pytest.raises(KeyError):
        LightGBMRegressor(weightCol="nope")._named_columns(LocalDataFrame([{"features": X, "label": y}]))
    same = LocalDataFrame([{"features": X, "label": y}])
    assert LightGBMRegressor()._named_columns(same) is same


def test_thresholds_pick_the_class_as_spark_does():
    # ProbabilisticClassificationModel.probability2prediction: arg max of probability / threshold, first maximum wins, a zero threshold
    # always wins unless its probability is zero too
    from synapseml_b200.lightgbm.estimators import LightGBMClassificationModel
    f = LightGBMClassificationModel.probability_to_prediction
    p = np.array([[0.6, 0.4], [0.3, 0.7], [0.5, 0.5]])
    assert f(p, [0.5, 0.5]).tolist() == [0.0, 1.0, 0.0]                       # a tie keeps the first class
    assert f(p, [0.9, 0.1]).tolist() == [1.0, 1.0, 1.0]
    assert f(p, [0.0, 1.0]).tolist() == [0.0, 0.0, 0.0]                       # +inf wins
    assert f(np.array([[0.0, 1.0], [0.2, 0.8]]), [0.0, 1.0]).tolist() == [1.0, 0.0]      # 0 / 0 never wins
    p3 = np.array([[0.2, 0.3, 0.5], [0.1, 0.1, 0.8]])
    assert f(p3, [0.1, 0.6, 1.0]).tolist() == [0.0, 0.0] and f(p3, [1.0, 1.0, 1.0]).tolist() == [2.0, 2.0]
    assert f(np.zeros((1, 3)), [0.0, 0.0, 0.0]).tolist() == [0.0]             # nothing compares greater: class 0
