"""LightGBMClassifier / LightGBMRegressor / LightGBMRanker: the estimator API of the reference with its parameter names
and defaults (LightGBMParams.scala), over a local stand-in for a Spark DataFrame.

Mirrors LightGBMBase.scala (train :36-65, trainOneDataBatch :396-447, calculateRowStatistics :509-550, getSampledRows
:724-749, executeTraining :569-606), LightGBMClassifier.scala:27-89, LightGBMRegressor.scala:39-92 and
LightGBMRanker.scala:26-121.  One process is one executor with one GPU; under torchrun every rank calls fit() with its
own partitions (rows of rank r precede rows of rank r+1 in the global order).
"""
import re

import numpy as np

from . import network_manager as nm
from .dataset import ReferenceDatasetUtils
from .booster import LightGBMBooster
from .model_text import TextModel
from .params import (CategoricalParams, ClassifierTrainParams, DartModeParams, DatasetParams, ExecutionParams, GeneralParams,
                     ObjectiveParams, QuantizedGradParams, RankerTrainParams, RegressorTrainParams, SeedParams,
                     dataset_creation_params)
from .partition_task import BulkPartitionTask, StreamingPartitionTask


class LocalDataFrame:
    """Stand-in for a Spark DataFrame: a list of partitions, each a dict column -> numpy array."""

    def __init__(self, partitions):
        self.partitions = list(partitions)

    @staticmethod
    def from_arrays(features, label, num_partitions=1, **cols):
        n = len(label)
        bounds = [n * i // num_partitions for i in range(num_partitions + 1)]
        parts = []
        for i in range(num_partitions):
            b, e = bounds[i], bounds[i + 1]
            p = {"features": features[b:e], "label": label[b:e]}
            for k, v in cols.items():
                if v is not None:
                    p[k] = v[b:e]
            parts.append(p)
        return LocalDataFrame(parts)

    def count(self):
        return sum(len(p["label"]) for p in self.partitions)


def _split_validation(df, vcol):
    train_parts, valid_parts = [], []
    for p in df.partitions:
        mask = np.asarray(p[vcol]).astype(bool)
        keep = ~mask
        train_parts.append({k: (v[keep] if k != vcol else v[keep]) for k, v in p.items() if k != vcol})
        if mask.any():
            valid_parts.append({k: v[mask] for k, v in p.items() if k != vcol})
    if not valid_parts:
        return LocalDataFrame(train_parts), None
    keys = valid_parts[0].keys()
    return LocalDataFrame(train_parts), {k: np.concatenate([vp[k] for vp in valid_parts]) for k in keys}


# name -> default.  None = "not set": the key is then omitted from the native string and the native default applies
# (LightGBMBase.scala:249-288 uses get(param), which ignores Spark-side defaults).
_DEFAULTS = dict(
    numIterations=100, learningRate=0.1, numLeaves=None, maxBin=None, binSampleCount=None, baggingFraction=None,
    posBaggingFraction=None, negBaggingFraction=None, baggingFreq=None, baggingSeed=None, earlyStoppingRound=0,
    improvementTolerance=0.0, featureFraction=None, featureFractionByNode=None, maxDepth=None, minSumHessianInLeaf=None,
    parallelism="data_parallel", topK=None, verbosity=-1, boostingType="gbdt", lambdaL1=None, lambdaL2=None, metric=None,
    minGainToSplit=None, maxDeltaStep=None, minDataPerBin=None, minDataInLeaf=None, topRate=None, otherRate=None,
    isEnableSparse=None, useMissing=None, zeroAsMissing=None, seed=None, deterministic=None, dataRandomSeed=None,
    passThroughArgs=None, numThreads=None, dataTransferMode="streaming", samplingMode="subset", samplingSubsetSize=1000000,
    microBatchSize=100, useSingleDatasetMode=True, maxStreamingOMPThreads=16, chunkSize=10000, matrixType="auto",
    labelCol="label", featuresCol="features", weightCol=None, initScoreCol=None, slotNames=None, referenceDataset=None,
    useQuantizedGrad=None, numGradQuantBins=None, stochasticRounding=None, isProvideTrainingMetric=None, delegate=None, validationIndicatorCol=None,
    defaultListenPort=nm.DEFAULT_LOCAL_LISTEN_PORT, timeout=1200.0, modelString=None,
    # categorical features (LightGBMParams.scala:286-296, 451-483): slot indexes and / or slot names (resolved against slotNames)
    categoricalSlotIndexes=None, categoricalSlotNames=None, minDataPerGroup=None, maxCatThreshold=None, catl2=None, catSmooth=None,
    maxCatToOnehot=None,
    monotoneConstraints=None, monotoneConstraintsMethod=None, monotonePenalty=None,      # LightGBMParams.scala (method "basic" is built)
    # DART (LightGBMDartParams) and the remaining seeds (LightGBMSeedParams)
    dropRate=0.1, maxDrop=50, skipDrop=0.5, xGBoostDartMode=False, uniformDrop=False,
    featureFractionSeed=None, extraSeed=None, dropSeed=None, objectiveSeed=None,
    # numBatches > 0: the rows are split into that many random batches which are trained one after the other, every batch continuing
    # from the model of the one before it (LightGBMBase.scala:39-64)
    numBatches=0,
    # model output columns (LightGBMPredictionParams): empty = not produced; startIteration: first iteration used for prediction
    leafPredictionCol="", featuresShapCol="", startIteration=0, predictDisableShapeCheck=False,
    # Spark's predictor columns: an empty name leaves the column out of transform's result (LightGBMClassifier.scala:129-146);
    # thresholds (ProbabilisticClassificationModel): one per class, the predicted class maximises probability / threshold
    predictionCol="prediction", rawPredictionCol="rawPrediction", probabilityCol="probability", thresholds=None,
    maxNumClasses=100,      # LightGBMClassifier: more label values than this is taken for a regression problem and refused
    # Spark-only controls, accepted and without effect here (one process per GPU, launched by torchrun: no Spark tasks to cap,
    # barrier or repartition, no driver socket)
    numTasks=0, useBarrierExecutionMode=False, repartitionByGroupingColumn=True, driverListenPort=0,
    maxBinByFeature=None,      # LightGBMParams.scala: sent with the TRAIN parameters only (LightGBMBase.scala:374-386 leaves it out of the
                               # dataset parameters), where the native side can only warn that the dataset is already built — as with lightgbmlib
    fobj=None,      # FObjTrait: an object with get_gradient(predictions [n][1], training_dataset) -> (grad f32[n], hess f32[n])
)
_SPARK_DEFAULTS = dict(maxBin=255, binSampleCount=200000, minDataInLeaf=20, numLeaves=31)   # values getXxx returns when unset


class LightGBMBase:
    _objective_default = "regression"

    def __init__(self, **kwargs):
        self._p = dict(_DEFAULTS)
        self._p["objective"] = self._objective_default
        self._p.update(self._extra_defaults())
        for k, v in kwargs.items():
            self._set(k, v)
        self.performance_measures = None

    def _extra_defaults(self):
        return {}

    def _set(self, k, v):
        if k not in self._p:
            raise AttributeError("Unknown param %s" % k)
        if k == "dataTransferMode" and v not in ("streaming", "bulk"):
            raise ValueError("dataTransferMode must be streaming or bulk")     # TrainingContext.scala:67-69
        self._p[k] = v
        return self

    def __getattr__(self, name):
        if name.startswith("set") and len(name) > 3:
            key = name[3].lower() + name[4:]
            if key == "executionMode":           # deprecated alias (LightGBMParams.scala:86-90)
                key = "dataTransferMode"
            return lambda v: self._set(key, v)
        if name.startswith("get") and len(name) > 3:
            key = name[3].lower() + name[4:]
            if key == "executionMode":
                key = "dataTransferMode"
            return lambda: self._p[key] if self._p.get(key) is not None else _SPARK_DEFAULTS.get(key)
        raise AttributeError(name)

    def get_categorical_indexes(self):
        """LightGBMBase.scala:167-198 (getCategoricalIndexes): categoricalSlotNames looked up in slotNames (a name that is not there
        resolves to -1 there too, which the native side then refuses), united with categoricalSlotIndexes.  (A LocalDataFrame carries
        no ML attribute metadata, so nominal attributes cannot mark columns here.)"""
        g = self._p
        names = list(g["categoricalSlotNames"] or [])
        slots = list(g["slotNames"] or [])
        if names and not slots:
            raise ValueError("categoricalSlotNames needs slotNames to resolve the names against")
        by_name = [slots.index(n) if n in slots else -1 for n in names]
        out = []
        for i in list(g["categoricalSlotIndexes"] or []) + by_name:
            if int(i) not in out:
                out.append(int(i))
        return out

    # ------------------------------------------------------------------ param groups (LightGBMBase.scala:240-372)
    def _general_params(self, num_tasks):
        g = self._p
        return GeneralParams(categoricalFeatures=self.get_categorical_indexes(), maxBinByFeature=list(g["maxBinByFeature"] or []), monotoneConstraints=list(g["monotoneConstraints"] or []),
                             monotoneConstraintsMethod=g["monotoneConstraintsMethod"], monotonePenalty=g["monotonePenalty"], parallelism=g["parallelism"], topK=g["topK"], numIterations=g["numIterations"], learningRate=g["learningRate"],
                             numLeaves=g["numLeaves"], maxBin=g["maxBin"], binSampleCount=g["binSampleCount"],
                             baggingFraction=g["baggingFraction"], posBaggingFraction=g["posBaggingFraction"],
                             negBaggingFraction=g["negBaggingFraction"], baggingFreq=g["baggingFreq"], baggingSeed=g["baggingSeed"],
                             earlyStoppingRound=g["earlyStoppingRound"], improvementTolerance=g["improvementTolerance"],
                             featureFraction=g["featureFraction"], featureFractionByNode=g["featureFractionByNode"], maxDepth=g["maxDepth"],
                             minSumHessianInLeaf=g["minSumHessianInLeaf"], numMachines=num_tasks, modelString=g["modelString"],
                             verbosity=g["verbosity"], boostingType=g["boostingType"], lambdaL1=g["lambdaL1"], lambdaL2=g["lambdaL2"],
                             metric=g["metric"], minGainToSplit=g["minGainToSplit"], maxDeltaStep=g["maxDeltaStep"],
                             minDataPerBin=g["minDataPerBin"], minDataInLeaf=g["minDataInLeaf"], topRate=g["topRate"], otherRate=g["otherRate"],
                             featureNames=list(g["slotNames"] or []))

    def _common_groups(self, num_tasks, num_tasks_per_exec):
        g = self._p
        exec_threads = g["numThreads"] if g["numThreads"] is not None else max(num_tasks_per_exec - 1, 0)
        return dict(
            passThroughArgs=g["passThroughArgs"], isProvideTrainingMetric=g["isProvideTrainingMetric"], delegate=g["delegate"],
            generalParams=self._general_params(num_tasks),
            datasetParams=DatasetParams(g["isEnableSparse"], g["useMissing"], g["zeroAsMissing"]),
            dartModeParams=DartModeParams(g["dropRate"], g["maxDrop"], g["skipDrop"], g["xGBoostDartMode"], g["uniformDrop"]),
            executionParams=ExecutionParams(g["chunkSize"], g["matrixType"], exec_threads, g["dataTransferMode"], g["samplingMode"],
                                            g["samplingSubsetSize"], g["microBatchSize"], g["useSingleDatasetMode"], g["maxStreamingOMPThreads"]),
            objectiveParams=ObjectiveParams(g["objective"], g["fobj"]),
            seedParams=SeedParams(seed=g["seed"], deterministic=g["deterministic"], dataRandomSeed=g["dataRandomSeed"],
                                  baggingSeed=g["baggingSeed"], featureFractionSeed=g["featureFractionSeed"], extraSeed=g["extraSeed"],
                                  dropSeed=g["dropSeed"], objectiveSeed=g["objectiveSeed"],
                                  boostingType=g["boostingType"], objective=g["objective"]),
            categoricalParams=CategoricalParams(g["minDataPerGroup"], g["maxCatThreshold"], g["catl2"], g["catSmooth"], g["maxCatToOnehot"]),
            quantizedParams=QuantizedGradParams(g["useQuantizedGrad"], g["numGradQuantBins"], g["stochasticRounding"]))

    def get_train_params(self, num_tasks, num_tasks_per_exec):
        raise NotImplementedError

    # ------------------------------------------------------------------ sampling (LightGBMBase.scala:724-749)
    def get_sampled_rows(self, df, total_num_rows):
        g = self._p
        bin_sample_count = self.getBinSampleCount()
        num_samples = min(bin_sample_count, total_num_rows)
        mode = g["samplingMode"]
        feats = [p["features"] for p in df.partitions if len(p["label"])]

        def head(k):      # df.limit(k) without materialising the whole frame
            out, need = [], k
            for f in feats:
                if need <= 0:
                    break
                part = f[:need]
                out.append(part.toarray() if hasattr(part, "toarray") else part)      # sparse rows are sampled as dense values
                need -= out[-1].shape[0]
            return out[0] if len(out) == 1 else np.concatenate(out)

        if mode == "fixed":       # df.limit(numSamples)
            return head(num_samples)
        seed = g["dataRandomSeed"] if g["dataRandomSeed"] is not None else 0
        fraction = 1.0 if total_num_rows < bin_sample_count else min(1.0, (bin_sample_count + 10000.0) / total_num_rows)
        rng = np.random.Generator(np.random.PCG64(seed))
        if mode == "global":      # sample(fraction, seed).limit(numSamples)
            allx = np.concatenate(feats)
            keep = rng.random(len(allx)) < fraction
            return allx[keep][:num_samples]
        # "subset" (default): limit(samplingSubsetSize).sample(fraction, seed), with samplingSubsetSize capped at numSamples (:733)
        subset = min(g["samplingSubsetSize"], num_samples)
        hd = head(subset)
        keep = rng.random(len(hd)) < fraction
        return hd[keep]

    def _sample_across_ranks(self, df, total_rows, counts, rank, dist):
        """The reference samples the WHOLE DataFrame on the driver (df.limit / df.sample run over all partitions).  With one process
        per GPU the rows of the other ranks are not in this process: in "fixed" mode (the first numSamples rows of the global order)
        every rank hands the part of its rows that falls inside that head to rank 0; the other modes sample rank 0's rows only."""
        if dist is None or self._p["samplingMode"] != "fixed":
            return self.get_sampled_rows(df, total_rows) if rank == 0 else None
        import torch
        num_samples = min(self.getBinSampleCount(), total_rows)
        take = []      # rows of every rank that fall inside the first num_samples rows of the global order
        before = 0
        for c in counts:
            take.append(int(min(max(num_samples - before, 0), c)))
            before += c
        k = take[rank]
        mine = np.ascontiguousarray(self.get_sampled_rows(df, k), dtype=np.float64) if k > 0 else None      # "fixed": the first k local rows
        if rank != 0:
            if k > 0:
                dist.send(torch.from_numpy(mine), dst=0)      # a plain buffer: no pickling of what can be gigabytes (C5: 200k x 2000 f64)
            return None
        ncol = mine.shape[1]
        parts = [mine]
        for r in range(1, len(counts)):
            if take[r] > 0:
                buf = torch.empty((take[r], ncol), dtype=torch.float64)
                dist.recv(buf, src=r)
                parts.append(buf.numpy())
        return parts[0] if len(parts) == 1 else np.concatenate(parts)

    def preprocess_data(self, df):
        return df

    # ------------------------------------------------------------------ train
    def fit(self, df):
        """LightGBMBase.train (LightGBMBase.scala:33-65): one data batch, or numBatches random batches trained in sequence, each
        continuing from the model string of the previous one."""
        nb = int(self._p.get("numBatches") or 0)
        if nb <= 0:
            return self._train_one_data_batch(df)
        # dataset.randomSplit(equal weights): every row draws its batch (Spark's split is seeded at random; here by `seed`)
        rng = np.random.Generator(np.random.PCG64(self._p["seed"] if self._p["seed"] is not None else 0))
        draws = [rng.integers(0, nb, len(p[self._p["labelCol"]])) for p in df.partitions]
        saved = self._p["modelString"]
        model = None
        self.batch_performance_measures = []
        try:
            for b in range(nb):
                parts = [{k: v[d == b] for k, v in p.items()} for p, d in zip(df.partitions, draws)]
                if model is not None:
                    self._p["modelString"] = model.getNativeModel()
                model = self._train_one_data_batch(LocalDataFrame(parts))
                self.batch_performance_measures.append(self.performance_measures)
        finally:
            self._p["modelString"] = saved
        return model

    def _named_columns(self, df):
        """The column-name params (labelCol, featuresCol, weightCol, initScoreCol, groupCol — LightGBMParams / Spark's HasLabelCol ...)
        pick the columns of the frame; the partition tasks below work on the fixed keys "label", "features", "weight", "initScore",
        "group".  A weight / init-score column is used when its param names it — or, as before, when the frame simply carries a
        column of the fixed name."""
        g = self._p
        names = {"label": g["labelCol"], "features": g["featuresCol"]}
        if g.get("weightCol"):
            names["weight"] = g["weightCol"]
        if g.get("initScoreCol"):
            names["initScore"] = g["initScoreCol"]
        if g.get("groupCol"):
            names["group"] = g["groupCol"]
        if all(k == v for k, v in names.items()):
            return df
        parts = []
        for p in df.partitions:
            q = dict(p)
            for key, col in names.items():
                if col != key:
                    if col not in p:
                        raise KeyError("column %r (%sCol) is not in the data" % (col, key))
                    q[key] = p[col]
            parts.append(q)
        return LocalDataFrame(parts)

    def _train_one_data_batch(self, df):
        """trainOneDataBatch: row statistics, bin sample -> reference dataset, then the partition task."""
        import time
        df = self._named_columns(df)
        import torch.distributed as dist
        g = self._p
        t_start = time.time()
        df = self.preprocess_data(df)
        validation_data = None
        vcol = g.get("validationIndicatorCol")
        if vcol:
            # LightGBMBase.scala: rows flagged by validationIndicatorCol are collected and broadcast; training uses the rest
            df, validation_data = _split_validation(df, vcol)
        dist_on = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
        rank = dist.get_rank() if dist_on else 0
        world = dist.get_world_size() if dist_on else 1
        num_tasks_per_exec = max(1, len(df.partitions))
        first = next((p for p in df.partitions if len(p["label"])), None)
        if first is None:
            raise ValueError("Empty dataset")
        num_cols = first["features"].shape[1]
        local_rows = df.count()
        total_rows = local_rows
        if dist_on:
            counts = [None] * world
            dist.all_gather_object(counts, local_rows)
            total_rows = int(sum(counts))
        train_params = self.add_custom_train_params(self.get_train_params(world, num_tasks_per_exec), df, dist if dist_on else None)
        ds_params = dataset_creation_params(self.getMaxBin(), self.getBinSampleCount(), self.getMinDataInLeaf(),
                                            train_params.executionParams.numThreads, categorical_indexes=self.get_categorical_indexes(),
                                            data_random_seed=g["dataRandomSeed"] if g["dataRandomSeed"] is not None else g["seed"])
        if g["passThroughArgs"]:
            ds_params = g["passThroughArgs"] + " " + ds_params
        measures = {}
        ctx = dict(train_params=train_params, dataset_params=ds_params, num_cols=num_cols, feature_names=g["slotNames"],
                   measures=measures, hist_timing=getattr(self, "_hist_timing", False), keep_scores=getattr(self, "_keep_scores", False),
                   # one pushing thread per partition, as Spark runs one task thread per partition on an executor
                   # (BasePartitionTask.scala:84-112); the native push is thread-safe for distinct tid
                   threaded_push=getattr(self, "_threaded_push", True), validation_data=validation_data)
        if g["dataTransferMode"] == "streaming":
            serialized = g["referenceDataset"]
            if serialized is None:
                t0 = time.time()
                payload = [None]
                sampled = self._sample_across_ranks(df, total_rows, counts if dist_on else None, rank, dist if dist_on else None)
                if rank == 0:     # the driver samples and finds the bin boundaries (K1), everyone else receives the bytes
                    payload[0] = ReferenceDatasetUtils.create_reference_dataset_from_sample(ds_params, g["featuresCol"], total_rows,
                                                                                            num_cols, sampled)
                if dist_on:
                    dist.broadcast_object_list(payload, src=0)
                serialized = payload[0]
                measures["samplingTime"] = time.time() - t0
                self._p["referenceDataset"] = serialized      # cached across fits (LightGBMBase.scala:427-431)
            ctx["serialized_reference"] = serialized
            task = StreamingPartitionTask(ctx)
        else:
            task = BulkPartitionTask(ctx)
        ctx["network_info"] = nm.get_global_network_info(g["defaultListenPort"]) if dist_on else None
        t0 = time.time()
        model_str = task.run(df.partitions)
        measures["trainingTime"] = time.time() - t0
        measures["totalTime"] = time.time() - t_start
        self.performance_measures = measures
        self.best_iteration = measures.get("bestIteration")
        if dist_on:
            payload = [model_str]
            dist.broadcast_object_list(payload, src=0)
            model_str = payload[0]
        model = self._make_model(model_str)
        model.best_iteration = self.best_iteration
        # getModel(...).setNumIterations(lightGBMBooster.bestIteration) (LightGBMClassifier.scala:80): after early stopping the model
        # predicts with bestIteration iterations — the 0-based index, i.e. one tree set fewer than saveToString kept; -1 otherwise
        model.numIterations = -1 if self.best_iteration is None else int(self.best_iteration)
        model.leafPredictionCol, model.featuresShapCol = g["leafPredictionCol"], g["featuresShapCol"]
        model.startIteration = g["startIteration"]
        model.predictDisableShapeCheck = bool(g["predictDisableShapeCheck"])
        model.predictionCol, model.rawPredictionCol, model.probabilityCol = g["predictionCol"], g["rawPredictionCol"], g["probabilityCol"]
        model.thresholds = None if g["thresholds"] is None else [float(t) for t in g["thresholds"]]
        return model

    def add_custom_train_params(self, train_params, df, dist):
        """LightGBMBase.addCustomTrainParams: estimator-specific additions that need the data (the classifier's numClass)."""
        return train_params

    def getPerformanceMeasures(self):
        return self.performance_measures

    def _make_model(self, model_str):
        raise NotImplementedError


class LightGBMModelBase:
    """LightGBMModelMethods / the *Model classes: holds the model string; scoring rebuilds a native booster from it on first use
    (LightGBMBooster.scala:41-49) and predicts on the GPU through LGBM_BoosterPredictForMat."""

    def __init__(self, model_str):
        self.model_str = model_str
        self._parsed = None                       # the text parsed in Python, on first use (tests and tools; scoring is native)
        m = re.search(r"^num_tree_per_iteration=(\d+)", model_str, flags=re.M)
        if m is None:
            raise ValueError("Model string has no num_tree_per_iteration line")
        self._trees_per_iteration = int(m.group(1))
        self._booster = None
        # LightGBMPredictionParams: optional output columns of transform, and the first iteration used for prediction
        self.leafPredictionCol, self.featuresShapCol, self.startIteration = "", "", 0
        self.predictDisableShapeCheck = False
        self.predictionCol, self.rawPredictionCol, self.probabilityCol, self.thresholds = "prediction", "rawPrediction", "probability", None
        self.numIterations = -1      # LightGBMPredictionParams.numIterations: iterations used for prediction, <= 0 = all from startIteration

    def setNumIterations(self, v):
        self.numIterations = int(v)
        return self

    @property
    def _model(self):
        if self._parsed is None:
            self._parsed = TextModel(self.model_str)
        return self._parsed

    def setPredictionCol(self, v):
        self.predictionCol = v
        return self

    def getLightGBMBooster(self):
        """LightGBMModelParams.getLightGBMBooster: the booster wrapper the model scores with."""
        return self._native()

    def setLeafPredictionCol(self, v):
        self.leafPredictionCol = v
        return self

    def setFeaturesShapCol(self, v):
        self.featuresShapCol = v
        return self

    def setStartIteration(self, v):
        self.startIteration = int(v)
        return self

    def _extra_columns(self, features, out):
        """The leaf-index and SHAP columns transform adds when their names are set (LightGBMClassifier.scala:151-160)."""
        if self.leafPredictionCol:
            out[self.leafPredictionCol] = self.predict_leaf(features)
        if self.featuresShapCol:
            out[self.featuresShapCol] = self.predict_contrib(features)
        return out

    def _predict(self, features, predict_type):
        # updateBoosterParamsBeforePredict (LightGBMModelMethods.scala:129-132): the model's startIteration / numIterations go to the booster
        return self._native().predict_for_mat(features, predict_type, start_iteration=self.startIteration, num_iteration=self.numIterations,
                                              disable_shape_check=self.predictDisableShapeCheck)

    def setPredictDisableShapeCheck(self, v):
        self.predictDisableShapeCheck = bool(v)
        return self

    def _native(self):
        if self._booster is None:
            self._booster = LightGBMBooster(model_str=self.model_str)
        return self._booster

    def __del__(self):
        try:
            if self._booster is not None:
                self._booster.free_native_memory()
        except Exception:  # noqa: BLE001
            pass

    def getNativeModel(self):
        return self.model_str

    def saveNativeModel(self, path):
        with open(path, "w") as f:
            f.write(self.model_str)

    def predict_raw(self, features):
        """Raw scores: [n], or [n, num_class] for a multiclass model."""
        raw = self._predict(features, LightGBMBooster.PREDICT_RAW)
        K = self._trees_per_iteration
        return raw if K == 1 else raw.reshape(-1, K)

    def predict_leaf(self, features):
        return self._predict(features, LightGBMBooster.PREDICT_LEAF).astype(np.int64)

    def predict_contrib(self, features):
        """featuresShapCol (LightGBMBooster.scala:418-427, predict type C_API_PREDICT_CONTRIB): per row, the TreeSHAP contribution of
        every feature followed by the expected value — [n, numFeatures + 1], or [n, numClasses * (numFeatures + 1)] for multiclass.
        A row's contributions sum to its raw score."""
        return self._predict(features, LightGBMBooster.PREDICT_CONTRIB)

    def getFeatureImportances(self, importance_type="split"):
        return self._native().get_feature_importances(importance_type)

    # ---- LightGBMModelMethods.scala:28-96
    def getFeatureShaps(self, features):
        """Local SHAP values of one instance (featuresShap): numClasses * (numFeatures + 1) values."""
        return self._native().predict_for_mat(np.asarray(features, dtype=np.float64).reshape(1, -1), LightGBMBooster.PREDICT_CONTRIB)[0].tolist()

    def getDenseFeatureShaps(self, features):
        return self.getFeatureShaps(features)

    def getSparseFeatureShaps(self, size, indices, values):
        row = np.zeros(size, dtype=np.float64)
        row[np.asarray(indices, dtype=np.int64)] = np.asarray(values, dtype=np.float64)
        return self.getFeatureShaps(row)

    def getBoosterBestIteration(self):
        """-1 unless early stopping was triggered (LightGBMBooster.scala:247)."""
        return -1 if getattr(self, "best_iteration", None) is None else int(self.best_iteration)

    def getBoosterNumTotalModel(self):
        return self._native().num_total_model()

    def getBoosterNumTotalIterations(self):
        return self._native().num_total_model() // max(1, self._trees_per_iteration)      # LightGBMBooster.scala:518

    def getBoosterNumFeatures(self):
        return self._native().num_features()

    def getBoosterNumClasses(self):
        return self._native().num_classes()

    @classmethod
    def loadNativeModelFromString(cls, model):
        """LightGBMClassificationModel / RegressionModel / RankerModel .loadNativeModelFromString."""
        return cls(model)

    @classmethod
    def loadNativeModelFromFile(cls, filename):
        with open(filename) as f:
            return cls(f.read())


class LightGBMClassificationModel(LightGBMModelBase):
    def setRawPredictionCol(self, v):
        self.rawPredictionCol = v
        return self

    def setProbabilityCol(self, v):
        self.probabilityCol = v
        return self

    def setThresholds(self, v):
        self.thresholds = None if v is None else [float(t) for t in v]
        return self

    def getActualNumClasses(self):
        """HasActualNumClasses (LightGBMClassifier.scala:93-98, 202): the booster's class count — 2 for a binary model."""
        K = self._trees_per_iteration
        return K if K > 1 else 2

    @property
    def numClasses(self):
        return self.getActualNumClasses()

    def transform(self, features):
        """LightGBMClassificationModel.transform (LightGBMClassifier.scala:118-163): raw scores and probabilities come from the native
        predictor; the prediction is the arg max of the raw scores, or — with thresholds — of probability / threshold (Spark's
        probability2prediction: a class with threshold 0 wins unless its probability is 0 too).  Columns with an empty name are left out."""
        K = self._trees_per_iteration
        if self.thresholds is not None and len(self.thresholds) != self.numClasses:
            raise ValueError("LightGBMClassificationModel.transform() called with non-matching numClasses and thresholds.length. numClasses=%d, "
                             "but thresholds has length %d" % (self.numClasses, len(self.thresholds)))
        raw = self.predict_raw(features)
        p = self._predict(features, LightGBMBooster.PREDICT_NORMAL)
        # binary: the native value is the positive-class score; expanded as the reference does (LightGBMBooster.scala:559-575)
        raw2 = raw if K > 1 else np.stack([-raw, raw], axis=1)
        prob2 = p.reshape(-1, K) if K > 1 else np.stack([1 - p, p], axis=1)
        out = {}
        if self.rawPredictionCol:
            out[self.rawPredictionCol] = raw2
        if self.probabilityCol:
            out[self.probabilityCol] = prob2
        if self.predictionCol:
            out[self.predictionCol] = (np.argmax(raw2, axis=1).astype(np.float64) if self.thresholds is None
                                       else self.probability_to_prediction(prob2, self.thresholds))
        return self._extra_columns(features, out)

    @staticmethod
    def probability_to_prediction(probability, thresholds):
        """Spark's ProbabilisticClassificationModel.probability2prediction with thresholds: the first class that maximises
        probability / threshold.  A zero threshold scales to +inf and wins — unless its probability is 0 as well (0 / 0 is NaN, which
        never compares greater); with nothing greater than -inf the answer stays class 0."""
        with np.errstate(divide="ignore", invalid="ignore"):
            scaled = np.asarray(probability, dtype=np.float64) / np.asarray(thresholds, dtype=np.float64)[None, :]
        return np.argmax(np.where(np.isnan(scaled), -np.inf, scaled), axis=1).astype(np.float64)


class LightGBMRegressionModel(LightGBMModelBase):
    def transform(self, features):
        # LightGBMRegressionModel.predict: score(features, raw = false): the objective's output transform applies (exp for
        # poisson / gamma / tweedie, identity otherwise)
        return self._extra_columns(features, {self.predictionCol: self._predict(features, LightGBMBooster.PREDICT_NORMAL)} if self.predictionCol else {})


class LightGBMRankerModel(LightGBMModelBase):
    def transform(self, features):
        return self._extra_columns(features, {self.predictionCol: self.predict_raw(features)} if self.predictionCol else {})


class LightGBMClassifier(LightGBMBase):
    _objective_default = "binary"          # LightGBMClassifier.scala:35

    def _extra_defaults(self):
        return dict(isUnbalance=False, boostFromAverage=True)

    def get_train_params(self, num_tasks, num_tasks_per_exec):
        return ClassifierTrainParams(isUnbalance=self._p["isUnbalance"], boostFromAverage=self._p["boostFromAverage"],
                                     numClass=getattr(self, "_num_class", 1), **self._common_groups(num_tasks, num_tasks_per_exec))

    def add_custom_train_params(self, train_params, df, dist):
        """LightGBMClassifier.addCustomTrainParams (LightGBMClassifier.scala:59-67): for anything but the binary objective the number
        of classes is inferred from the data — Spark's getNumClasses without label metadata: the largest label + 1."""
        if self._p["objective"] == "binary":
            return train_params
        local = max((int(np.max(p["label"])) for p in df.partitions if len(p["label"])), default=0)
        if dist is not None:
            allmax = [None] * dist.get_world_size()
            dist.all_gather_object(allmax, local)
            local = max(allmax)
        self._num_class = local + 1
        if self._num_class > int(self._p["maxNumClasses"]):      # getNumClasses(dataset, maxNumClasses): a label column like that is not categorical
            raise ValueError("Classifier inferred %d classes from the label column, but at most %d are allowed (maxNumClasses)"
                             % (self._num_class, int(self._p["maxNumClasses"])))
        train_params.numClass = self._num_class
        return train_params

    def _make_model(self, model_str):
        return LightGBMClassificationModel(model_str)


class LightGBMRegressor(LightGBMBase):
    _objective_default = "regression"      # LightGBMRegressor.scala:47

    def _extra_defaults(self):
        return dict(alpha=0.9, tweedieVariancePower=1.5, boostFromAverage=True)

    def get_train_params(self, num_tasks, num_tasks_per_exec):
        return RegressorTrainParams(alpha=self._p["alpha"], tweedieVariancePower=self._p["tweedieVariancePower"],
                                    boostFromAverage=self._p["boostFromAverage"], **self._common_groups(num_tasks, num_tasks_per_exec))

    def _make_model(self, model_str):
        return LightGBMRegressionModel(model_str)


class LightGBMRanker(LightGBMBase):
    _objective_default = "lambdarank"      # LightGBMRanker.scala:34

    def _extra_defaults(self):
        return dict(maxPosition=20, labelGain=[], evalAt=[1, 2, 3, 4, 5], groupCol="group")

    def get_train_params(self, num_tasks, num_tasks_per_exec):
        return RankerTrainParams(maxPosition=self._p["maxPosition"], labelGain=list(self._p["labelGain"]), evalAt=list(self._p["evalAt"]),
                                 **self._common_groups(num_tasks, num_tasks_per_exec))

    def preprocess_data(self, df):
        """sortWithinPartitions(groupCol) (LightGBMRanker.scala:88-90); groups must not straddle partitions (:94-120)."""
        parts = []
        for p in df.partitions:
            if "group" not in p:
                raise ValueError("LightGBMRanker needs a group column")
            g = np.asarray(p["group"])
            if g.dtype.kind not in "iuUSO":
                raise ValueError("group column must be int, long or string")
            order = np.argsort(g, kind="stable")
            parts.append({k: (np.asarray(v)[order] if k != "_gid" else v) for k, v in p.items()})
        return LocalDataFrame(parts)

    def _make_model(self, model_str):
        return LightGBMRankerModel(model_str)
