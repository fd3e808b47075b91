"""Per-executor partition tasks: marshal rows into the native dataset, then train.

Mirror of BasePartitionTask.scala:118-462 (template: initialize -> preparePartitionData -> network init ->
finalizeDatasetAndTrain -> cleanup), StreamingPartitionTask.scala:17-365 (micro-batches of `microBatchSize` rows pushed
with LGBM_DatasetPushRowsWithMetadata at the partition's row offset, :173-236) and BulkPartitionTask.scala:14-137 /
DatasetAggregator.scala:405-429 (whole executor matrix through LGBM_DatasetCreateFromMat + SetField).

One executor == one process == one GPU.  `partitions` is the list of this executor's partitions, each a dict with
"features" (float64 [n, F]), "label", and optionally "weight", "initScore", "group".
"""
import ctypes as C
import threading
import time

import numpy as np

from . import network_manager as nm
from . import train_utils
from .dataset import GroupIdManager, LightGBMDataset, ReferenceDatasetUtils
from .native import lightgbmlib, validate


def _as_f32(a):
    # labels and weights travel as float32 (StreamingPartitionTask.scala:332-336)
    return np.ascontiguousarray(a, dtype=np.float32)


class BasePartitionTask:
    def __init__(self, ctx):
        self.ctx = ctx

    def run(self, partitions):
        ctx = self.ctx
        partitions = [p for p in partitions if len(p["label"]) > 0]      # empty partitions short-circuit (:134-136)
        measures = ctx.setdefault("measures", {})
        t0 = time.time()
        dataset = self.prepare_dataset(partitions)
        measures["datasetCreationTime"] = time.time() - t0      # LightGBMPerformance.scala: the executor's dataset phase
        info = ctx.get("network_info")
        t0 = time.time()
        if info is not None:
            nm.init_lightgbm_network(info)
        measures["networkInitTime"] = time.time() - t0
        booster = None
        valid_ds = None
        try:
            dataset.validate_dataset()
            if ctx.get("feature_names"):
                dataset.set_feature_names(ctx["feature_names"], ctx["num_cols"])
            t0 = time.time()
            booster = train_utils.create_booster(ctx["train_params"], dataset)
            measures["boosterCreationTime"] = time.time() - t0
            measures["boosterSetupMs"] = booster.get_profile()["setup_ms"]
            if ctx.get("validation_data") is not None:
                valid_ds = self.generate_validation_dataset(ctx["validation_data"], dataset)
                booster.add_valid_data(valid_ds)
            if ctx.get("hist_timing"):
                booster.set_hist_timing(True)
            best_iteration = train_utils.execute_training_iterations(ctx["train_params"], booster, measures, valid_ds is not None)
            measures["bestIteration"] = best_iteration
            measures.update(booster.stats())
            # only the main worker of rank 0 returns the model (BasePartitionTask.getPartitionTaskResult :450-461)
            t0 = time.time()
            # saveToString(bestIterationResult): iterations up to and including the best one (LightGBMBooster.scala:272-278)
            upper = None if best_iteration is None else best_iteration + 1
            model_str = booster.save_to_string(upper) if (info is None or info.rank == 0) else None
            measures["saveModelTime"] = time.time() - t0
            if ctx.get("keep_scores"):
                measures["train_scores"] = booster.inner_predict(0)
            return model_str
        finally:
            t0 = time.time()
            if booster is not None:
                booster.free_native_memory()
            if valid_ds is not None:
                valid_ds.close()
            dataset.close()
            if info is not None and info.world > 1:
                nm.free_lightgbm_network()
            measures["cleanupTime"] = time.time() - t0      # freeing the booster's and the dataset's device / pinned memory


class StreamingPartitionTask(BasePartitionTask):
    def generate_validation_dataset(self, vdata, train_dataset):
        """generateOptValidationDataset (StreamingPartitionTask.scala:154-171): a dataset with the training bin mappers, filled from
        the broadcast validation rows in one go and finished."""
        ctx = self.ctx
        n = len(vdata["label"])
        has_w, has_i, has_g = "weight" in vdata, "initScore" in vdata, "group" in vdata
        handle = ReferenceDatasetUtils.get_initialized_reference_dataset(ctx["serialized_reference"], n, 1, ctx["dataset_params"],
                                                                         has_w, has_i, has_g, 1)
        mb = int(ctx["train_params"].executionParams.microBatchSize)
        self.insert_rows_into_training_dataset(handle, vdata, 0, mb, 0, GroupIdManager(), has_w, has_i, has_g)
        validate(lightgbmlib().LGBM_DatasetMarkFinished(handle), "Dataset mark finished")
        return LightGBMDataset(handle)

    def prepare_dataset(self, partitions):
        ctx = self.ctx
        n_rows = sum(len(p["label"]) for p in partitions)
        has_w = any("weight" in p for p in partitions)
        has_i = any("initScore" in p for p in partitions)
        has_g = any("group" in p for p in partitions)
        n_threads = max(1, len(partitions))
        handle = ReferenceDatasetUtils.get_initialized_reference_dataset(ctx["serialized_reference"], n_rows, 1, ctx["dataset_params"],
                                                                         has_w, has_i, has_g, n_threads)
        # row offset of each partition within the executor's dataset (BasePartitionTask.scala:92-112)
        offsets = np.concatenate([[0], np.cumsum([len(p["label"]) for p in partitions])])
        group_ids = GroupIdManager()
        mb = int(ctx["train_params"].executionParams.microBatchSize)
        errors = []

        def push_partition(tid, part, offset):
            try:
                self.insert_rows_into_training_dataset(handle, part, int(offset), mb, tid, group_ids, has_w, has_i, has_g)
            except Exception as e:  # noqa: BLE001
                errors.append(e)

        if ctx.get("threaded_push", False) and len(partitions) > 1:
            if has_g:   # ids must be assigned in row order
                for tid, p in enumerate(partitions):
                    p["_gid"] = np.array([group_ids.get_unique_index_for_id(v) for v in p["group"].tolist()], dtype=np.int32)
            threads = [threading.Thread(target=push_partition, args=(tid, p, offsets[tid])) for tid, p in enumerate(partitions)]
            [t.start() for t in threads]
            [t.join() for t in threads]
        else:
            for tid, p in enumerate(partitions):
                push_partition(tid, p, offsets[tid])
        if errors:
            raise errors[0]
        validate(lightgbmlib().LGBM_DatasetMarkFinished(handle), "Dataset mark finished")
        return LightGBMDataset(handle)

    @staticmethod
    def insert_rows_into_training_dataset(handle, part, start_index, micro_batch_size, tid, group_ids, has_w, has_i, has_g):
        """pushDenseMicroBatches (StreamingPartitionTask.scala:203-236): row-major double[mb * numCols] + metadata arrays."""
        lib = lightgbmlib()
        X = part["features"]
        n, ncol = X.shape
        label = _as_f32(part["label"])
        weight = _as_f32(part["weight"]) if has_w else None
        init = np.ascontiguousarray(part["initScore"], dtype=np.float64) if has_i else None
        if has_g:
            gid = part.get("_gid")
            if gid is None:
                gid = np.array([group_ids.get_unique_index_for_id(v) for v in np.asarray(part["group"]).tolist()], dtype=np.int32)
        if hasattr(X, "indptr"):      # scipy CSR partition: pushSparseMicroBatches (StreamingPartitionTask.scala:239-277)
            X = X.tocsr()
            indptr = np.ascontiguousarray(X.indptr, dtype=np.int64)
            indices = np.ascontiguousarray(X.indices, dtype=np.int32)
            values = np.ascontiguousarray(X.data, dtype=np.float64)
            for s in range(0, n, micro_batch_size):
                e = min(n, s + micro_batch_size)
                lo, hi = int(indptr[s]), int(indptr[e])
                ip = indptr[s:e + 1] - lo
                validate(lib.LGBM_DatasetPushRowsByCSRWithMetadata(
                    handle, ip.ctypes.data_as(C.c_void_p), C.c_int(3), indices[lo:hi].ctypes.data_as(C.POINTER(C.c_int32)),
                    values[lo:hi].ctypes.data_as(C.c_void_p), C.c_int(1), C.c_int64(e - s + 1), C.c_int64(hi - lo), C.c_int64(start_index + s),
                    label[s:e].ctypes.data_as(C.POINTER(C.c_float)),
                    weight[s:e].ctypes.data_as(C.POINTER(C.c_float)) if has_w else None,
                    init[s:e].ctypes.data_as(C.POINTER(C.c_double)) if has_i else None,
                    gid[s:e].ctypes.data_as(C.POINTER(C.c_int32)) if has_g else None,
                    C.c_int32(tid)), "Dataset push CSR micro-batch")
            return
        for s in range(0, n, micro_batch_size):
            e = min(n, s + micro_batch_size)
            xb = np.ascontiguousarray(X[s:e], dtype=np.float64)
            validate(lib.LGBM_DatasetPushRowsWithMetadata(
                handle, xb.ctypes.data_as(C.c_void_p), C.c_int(1), C.c_int32(e - s), C.c_int32(ncol), C.c_int32(start_index + s),
                label[s:e].ctypes.data_as(C.POINTER(C.c_float)),
                weight[s:e].ctypes.data_as(C.POINTER(C.c_float)) if has_w else None,
                init[s:e].ctypes.data_as(C.POINTER(C.c_double)) if has_i else None,
                gid[s:e].ctypes.data_as(C.POINTER(C.c_int32)) if has_g else None,
                C.c_int32(tid)), "Dataset push micro-batch")


class BulkPartitionTask(BasePartitionTask):
    def generate_validation_dataset(self, vdata, train_dataset):
        """BulkPartitionTask.getValidationDatasetInternal (BulkPartitionTask.scala): the broadcast validation rows as one dense matrix,
        binned with the training dataset as reference (LGBM_DatasetCreateFromMat(..., reference = train))."""
        lib = lightgbmlib()
        X = np.ascontiguousarray(vdata["features"].toarray() if hasattr(vdata["features"], "toarray") else vdata["features"], dtype=np.float64)
        handle = C.c_void_p()
        validate(lib.LGBM_DatasetCreateFromMat(X.ctypes.data_as(C.c_void_p), C.c_int(1), C.c_int32(X.shape[0]), C.c_int32(X.shape[1]),
                                               C.c_int(1), self.ctx["dataset_params"].encode(), train_dataset.dataset_ptr, C.byref(handle)),
                 "Dataset create")
        ds = LightGBMDataset(handle)
        ds.add_float_field(vdata["label"], "label")
        if "weight" in vdata:
            ds.add_float_field(vdata["weight"], "weight")
        if "initScore" in vdata:
            ds.add_double_field(vdata["initScore"], "init_score")
        if "group" in vdata:
            ds.add_group_column(np.asarray(vdata["group"]).tolist())
        return ds

    def prepare_dataset(self, partitions):
        """DenseAggregatedColumns.generateDataset (DatasetAggregator.scala:405-429): one LGBM_DatasetCreateFromMat, then
        label / weight / init_score / group through LGBM_DatasetSetField (LightGBMDataset.scala:110-176)."""
        ctx = self.ctx
        lib = lightgbmlib()
        handle = C.c_void_p()
        if any(hasattr(p["features"], "indptr") for p in partitions):
            # SparseAggregatedColumns.generateDataset (DatasetAggregator.scala): one CSR block for the executor
            import scipy.sparse as sp
            M = sp.vstack([sp.csr_matrix(p["features"]) for p in partitions]).tocsr()
            indptr = np.ascontiguousarray(M.indptr, dtype=np.int64)
            indices = np.ascontiguousarray(M.indices, dtype=np.int32)
            values = np.ascontiguousarray(M.data, dtype=np.float64)
            validate(lib.LGBM_DatasetCreateFromCSR(indptr.ctypes.data_as(C.c_void_p), C.c_int(3), indices.ctypes.data_as(C.POINTER(C.c_int32)),
                                                   values.ctypes.data_as(C.c_void_p), C.c_int(1), C.c_int64(len(indptr)), C.c_int64(len(values)),
                                                   C.c_int64(M.shape[1]), ctx["dataset_params"].encode(), None, C.byref(handle)), "Dataset create")
        else:
            X = np.ascontiguousarray(np.concatenate([p["features"] for p in partitions]), dtype=np.float64)
            validate(lib.LGBM_DatasetCreateFromMat(X.ctypes.data_as(C.c_void_p), C.c_int(1), C.c_int32(X.shape[0]), C.c_int32(X.shape[1]),
                                                   C.c_int(1), ctx["dataset_params"].encode(), None, C.byref(handle)), "Dataset create")
        ds = LightGBMDataset(handle)
        ds.add_float_field(np.concatenate([p["label"] for p in partitions]), "label")
        if any("weight" in p for p in partitions):
            ds.add_float_field(np.concatenate([p["weight"] for p in partitions]), "weight")
        if any("initScore" in p for p in partitions):
            ds.add_double_field(np.concatenate([p["initScore"] for p in partitions]), "init_score")
        if any("group" in p for p in partitions):
            ds.add_group_column(np.concatenate([np.asarray(p["group"]) for p in partitions]).tolist())
        return ds
