"""Parameter groups and the native parameter string.

Mirror of lightgbm/src/main/scala/com/microsoft/azure/synapse/ml/lightgbm/params/BaseTrainParams.scala (groups and their
append order, :25-40 and :99-137), ClassifierTrainParams.scala:28-33, RegressorTrainParams.scala:24-29,
RankerTrainParams.scala:24-29 and core/.../core/utils/ParamsStringBuilder.scala:100-175 ("append if not there": the
first writer of a key wins, which is how passThroughArgs overrides explicit setters).
"""
import re
from dataclasses import dataclass, field
from typing import List, Optional


def _fmt(v):
    """Value formatting as Scala's string interpolation prints it (true/false, Double.toString)."""
    if isinstance(v, bool):
        return "true" if v else "false"
    if isinstance(v, float):
        if v == int(v) and abs(v) < 1e7:
            return "%.1f" % v
        a = abs(v)
        if a != 0 and (a < 1e-3 or a >= 1e7):      # JVM Double.toString switches to E-notation here
            m, e = ("%.16e" % v).split("e")
            m = m.rstrip("0")
            if m.endswith("."):
                m += "0"
            return "%sE%d" % (m, int(e))
        return repr(v)
    return str(v)


class ParamsStringBuilder:
    def __init__(self, prefix="", delimiter="="):
        self.prefix, self.delimiter = prefix, delimiter
        self.sb = ""

    def append(self, s):
        if s:
            if self.sb:
                self.sb += " "
            self.sb += s
        return self

    def append_param_value_if_not_there(self, option_long, value):
        if value is not None and re.search(re.escape(self.prefix + option_long) + r"[ =]", self.sb) is None:
            self.append("%s%s%s%s" % (self.prefix, option_long, self.delimiter, _fmt(value)))
        return self

    def append_param_flag_if_not_there(self, name):
        if re.search(re.escape(self.prefix + name), self.sb) is None:
            self.append(self.prefix + name)
        return self

    def append_param_list_if_not_there(self, name, values):
        if values is not None and len(values) > 0 and re.search(re.escape(self.prefix + name), self.sb) is None:
            self.append("%s%s%s%s" % (self.prefix, name, self.delimiter, ",".join(_fmt(v) for v in values)))
        return self

    def append_param_group(self, group, condition=True):
        if condition:
            group.append_params(self)
        return self

    def result(self):
        return self.sb


@dataclass
class GeneralParams:
    parallelism: str = "data_parallel"
    topK: Optional[int] = None
    numIterations: int = 100
    learningRate: float = 0.1
    numLeaves: Optional[int] = None
    maxBin: Optional[int] = None
    binSampleCount: Optional[int] = None
    baggingFraction: Optional[float] = None
    posBaggingFraction: Optional[float] = None
    negBaggingFraction: Optional[float] = None
    baggingFreq: Optional[int] = None
    baggingSeed: Optional[int] = None
    earlyStoppingRound: int = 0
    improvementTolerance: float = 0.0
    featureFraction: Optional[float] = None
    featureFractionByNode: Optional[float] = None
    maxDepth: Optional[int] = None
    minSumHessianInLeaf: Optional[float] = None
    numMachines: int = 1
    modelString: Optional[str] = None
    categoricalFeatures: List[int] = field(default_factory=list)
    verbosity: int = -1
    boostingType: str = "gbdt"
    lambdaL1: Optional[float] = None
    lambdaL2: Optional[float] = None
    metric: Optional[str] = None
    minGainToSplit: Optional[float] = None
    maxDeltaStep: Optional[float] = None
    maxBinByFeature: List[int] = field(default_factory=list)
    minDataPerBin: Optional[int] = None
    minDataInLeaf: Optional[int] = None
    topRate: Optional[float] = None
    otherRate: Optional[float] = None
    monotoneConstraints: List[int] = field(default_factory=list)
    monotoneConstraintsMethod: Optional[str] = None
    monotonePenalty: Optional[float] = None
    featureNames: List[str] = field(default_factory=list)

    def append_params(self, sb):
        a = sb.append_param_value_if_not_there
        a("is_pre_partition", "True")
        a("boosting_type", self.boostingType)
        a("tree_learner", self.parallelism)
        a("top_k", self.topK)
        a("num_leaves", self.numLeaves)
        a("max_bin", self.maxBin)
        a("min_data_per_bin", self.minDataPerBin)
        a("bagging_fraction", self.baggingFraction)
        a("pos_bagging_fraction", self.posBaggingFraction)
        a("neg_bagging_fraction", self.negBaggingFraction)
        a("bagging_freq", self.baggingFreq)
        a("feature_fraction", self.featureFraction)
        a("max_depth", self.maxDepth)
        a("min_sum_hessian_in_leaf", self.minSumHessianInLeaf)
        a("lambda_l1", self.lambdaL1)
        a("lambda_l2", self.lambdaL2)
        a("metric", self.metric)
        a("min_gain_to_split", self.minGainToSplit)
        a("max_delta_step", self.maxDeltaStep)
        a("min_data_in_leaf", self.minDataInLeaf)
        a("num_iterations", self.numIterations)
        a("learning_rate", float(self.learningRate))
        a("num_machines", self.numMachines)
        a("verbosity", self.verbosity)
        a("early_stopping_round", self.earlyStoppingRound)
        a("learning_rate", float(self.learningRate))
        a("min_data_per_bin", self.minDataPerBin)
        sb.append_param_list_if_not_there("categorical_feature", self.categoricalFeatures)
        sb.append_param_list_if_not_there("max_bin_by_feature", self.maxBinByFeature)
        a("top_rate", self.topRate)
        a("other_rate", self.otherRate)
        sb.append_param_list_if_not_there("monotone_constraints", self.monotoneConstraints)
        a("monotone_constraints_method", self.monotoneConstraintsMethod)
        a("monotone_penalty", self.monotonePenalty)
        a("learning_rate", float(self.learningRate))


@dataclass
class DatasetParams:
    isEnableSparse: Optional[bool] = None
    useMissing: Optional[bool] = None
    zeroAsMissing: Optional[bool] = None

    def append_params(self, sb):
        sb.append_param_value_if_not_there("is_enable_sparse", self.isEnableSparse)
        sb.append_param_value_if_not_there("use_missing", self.useMissing)
        sb.append_param_value_if_not_there("zero_as_missing", self.zeroAsMissing)


@dataclass
class DartModeParams:
    dropRate: float = 0.1
    maxDrop: int = 50
    skipDrop: float = 0.5
    xgboostDartMode: bool = False
    uniformDrop: bool = False

    def append_params(self, sb):
        a = sb.append_param_value_if_not_there
        a("drop_rate", float(self.dropRate)); a("max_drop", self.maxDrop); a("skip_drop", float(self.skipDrop))
        a("xgboost_dart_mode", self.xgboostDartMode); a("uniform_drop", self.uniformDrop)


@dataclass
class ExecutionParams:
    chunkSize: int = 10000
    matrixType: str = "auto"
    numThreads: int = 0
    dataTransferMode: str = "streaming"
    samplingMode: str = "subset"
    samplingSetSize: int = 1000000
    microBatchSize: int = 100
    useSingleDatasetMode: bool = True
    maxStreamingOMPThreads: int = 16

    def append_params(self, sb):
        sb.append_param_value_if_not_there("num_threads", self.numThreads)


@dataclass
class ObjectiveParams:
    objective: str = "regression"
    fobj: Optional[object] = None

    def append_params(self, sb):
        sb.append_param_value_if_not_there("objective", self.objective if self.fobj is None else "custom")


@dataclass
class SeedParams:
    seed: Optional[int] = None
    deterministic: Optional[bool] = None
    baggingSeed: Optional[int] = None
    featureFractionSeed: Optional[int] = None
    extraSeed: Optional[int] = None
    dropSeed: Optional[int] = None
    dataRandomSeed: Optional[int] = None
    objectiveSeed: Optional[int] = None
    boostingType: str = "gbdt"
    objective: str = "regression"

    def append_params(self, sb):
        a = sb.append_param_value_if_not_there
        a("seed", self.seed); a("deterministic", self.deterministic); a("bagging_seed", self.baggingSeed)
        a("feature_fraction_seed", self.featureFractionSeed); a("extra_seed", self.extraSeed)
        a("data_random_seed", self.dataRandomSeed)
        a("drop_seed", self.dropSeed if self.boostingType == "dart" else None)
        a("objective_seed", self.objectiveSeed if self.objective == "rank_xendcg" else None)


@dataclass
class CategoricalParams:
    minDataPerGroup: Optional[int] = None
    maxCatThreshold: Optional[int] = None
    catl2: Optional[float] = None
    catSmooth: Optional[float] = None
    maxCatToOneHot: Optional[int] = None

    def append_params(self, sb):
        a = sb.append_param_value_if_not_there
        a("min_data_per_group", self.minDataPerGroup); a("max_cat_threshold", self.maxCatThreshold); a("cat_l2", self.catl2)
        a("cat_smooth", self.catSmooth); a("max_cat_to_onehot", self.maxCatToOneHot)


@dataclass
class QuantizedGradParams:
    """NEW in this project (the reference has no such param): useQuantizedGrad -> use_quantized_grad etc."""
    useQuantizedGrad: Optional[bool] = None
    numGradQuantBins: Optional[int] = None
    stochasticRounding: Optional[bool] = None

    def append_params(self, sb):
        a = sb.append_param_value_if_not_there
        a("use_quantized_grad", self.useQuantizedGrad); a("num_grad_quant_bins", self.numGradQuantBins)
        a("stochastic_rounding", self.stochasticRounding)


@dataclass
class BaseTrainParams:
    passThroughArgs: Optional[str] = None
    isProvideTrainingMetric: Optional[bool] = None
    delegate: Optional[object] = None
    generalParams: GeneralParams = field(default_factory=GeneralParams)
    datasetParams: DatasetParams = field(default_factory=DatasetParams)
    dartModeParams: DartModeParams = field(default_factory=DartModeParams)
    executionParams: ExecutionParams = field(default_factory=ExecutionParams)
    objectiveParams: ObjectiveParams = field(default_factory=ObjectiveParams)
    seedParams: SeedParams = field(default_factory=SeedParams)
    categoricalParams: CategoricalParams = field(default_factory=CategoricalParams)
    quantizedParams: QuantizedGradParams = field(default_factory=QuantizedGradParams)

    def get_num_class(self):
        return 1

    def append_specialized_params(self, sb):
        return sb

    def to_string(self):
        sb = ParamsStringBuilder(prefix="", delimiter="=")
        sb.append(self.passThroughArgs or "")
        sb.append_param_value_if_not_there("is_pre_partition", "True")
        sb.append_param_group(self.generalParams)
        sb.append_param_group(self.datasetParams)
        sb.append_param_group(self.dartModeParams, self.generalParams.boostingType == "dart")
        sb.append_param_group(self.objectiveParams)
        sb.append_param_group(self.executionParams)
        sb.append_param_group(self.seedParams)
        sb.append_param_group(self.categoricalParams)
        sb.append_param_group(self.quantizedParams)
        self.append_specialized_params(sb)
        return sb.result()

    __str__ = to_string


@dataclass
class ClassifierTrainParams(BaseTrainParams):
    isUnbalance: bool = False
    boostFromAverage: bool = True
    numClass: int = 1

    def get_num_class(self):
        return self.numClass

    def append_specialized_params(self, sb):
        is_binary = self.objectiveParams.objective == "binary"
        sb.append_param_value_if_not_there("num_class", self.numClass if not is_binary else None)
        sb.append_param_value_if_not_there("is_unbalance", self.isUnbalance if is_binary else None)
        sb.append_param_value_if_not_there("boost_from_average", self.boostFromAverage)
        return sb


@dataclass
class RegressorTrainParams(BaseTrainParams):
    alpha: float = 0.9
    tweedieVariancePower: float = 1.5
    boostFromAverage: bool = True

    def append_specialized_params(self, sb):
        sb.append_param_value_if_not_there("alpha", float(self.alpha))
        sb.append_param_value_if_not_there("tweedie_variance_power", float(self.tweedieVariancePower))
        sb.append_param_value_if_not_there("boost_from_average", self.boostFromAverage)
        return sb


@dataclass
class RankerTrainParams(BaseTrainParams):
    maxPosition: int = 20
    labelGain: List[float] = field(default_factory=list)
    evalAt: List[int] = field(default_factory=lambda: [1, 2, 3, 4, 5])

    def append_specialized_params(self, sb):
        sb.append_param_list_if_not_there("label_gain", [float(x) for x in self.labelGain])
        sb.append_param_list_if_not_there("eval_at", self.evalAt)
        sb.append_param_value_if_not_there("max_position", self.maxPosition)
        return sb


def dataset_creation_params(max_bin, bin_sample_count, min_data_in_leaf, num_threads, categorical_indexes=(), data_random_seed=None):
    """LightGBMBase.getDatasetCreationParams (LightGBMBase.scala:374-386)."""
    sb = ParamsStringBuilder(prefix="", delimiter="=")
    sb.append_param_value_if_not_there("is_pre_partition", "True")
    sb.append_param_value_if_not_there("max_bin", max_bin)
    sb.append_param_value_if_not_there("bin_construct_sample_cnt", bin_sample_count)
    sb.append_param_value_if_not_there("min_data_in_leaf", min_data_in_leaf)
    sb.append_param_value_if_not_there("num_threads", num_threads)
    sb.append_param_list_if_not_there("categorical_feature", list(categorical_indexes))
    sb.append_param_value_if_not_there("data_random_seed", data_random_seed)
    return sb.result()
