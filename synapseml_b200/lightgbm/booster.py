"""LightGBMBooster: mirror of lightgbm/src/main/scala/com/microsoft/azure/synapse/ml/lightgbm/booster/LightGBMBooster.scala
(training subset: create :241, updateOneIteration :355-365, resetParameter :316-323, saveToString :272-278,
getPredict :331-350, free :155), its scoring side (model string constructor :41-49, score / predictLeaf :407-427, 520-557) and
TrainUtils.createBooster (TrainUtils.scala:16-29)."""
import ctypes as C

import numpy as np

from .native import lightgbmlib, validate


class LightGBMBooster:
    def __init__(self, training_dataset=None, parameters=None, model_str=None):
        self.booster_handle = None
        self.model_str = model_str
        self.is_finished = False
        if training_dataset is not None:
            h = C.c_void_p()
            validate(lightgbmlib().LGBM_BoosterCreate(training_dataset.dataset_ptr, parameters.encode(), C.byref(h)), "Booster")
            self.booster_handle = h
            self.train_dataset = training_dataset
            self._train_rows = training_dataset.num_data()
        elif model_str is not None:
            # LightGBMBooster(modelStr): the booster is rebuilt from the model string where it is needed (:41-49)
            h, n_iter = C.c_void_p(), C.c_int()
            validate(lightgbmlib().LGBM_BoosterLoadModelFromString(model_str.encode(), C.byref(n_iter), C.byref(h)), "Booster LoadFromString")
            self.booster_handle = h
            self.num_iterations = n_iter.value

    # predict_type constants of the C API (C_API_PREDICT_NORMAL / RAW_SCORE / LEAF_INDEX)
    PREDICT_NORMAL, PREDICT_RAW, PREDICT_LEAF, PREDICT_CONTRIB = 0, 1, 2, 3

    def predict_for_mat(self, features, predict_type=0, start_iteration=0, num_iteration=-1, disable_shape_check=False):
        """Batch scoring on the GPU: one LGBM_BoosterPredictForMat call for all rows.  The parameter string is the reference's
        (LightGBMBooster.scala:526,548): max_bin=255 predict_disable_shape_check=<bool>."""
        lib = lightgbmlib()
        x = np.ascontiguousarray(features, dtype=np.float64)
        if x.ndim == 1:
            x = x.reshape(1, -1)
        n, ncol = x.shape
        need = C.c_int64()
        validate(lib.LGBM_BoosterCalcNumPredict(self.booster_handle, C.c_int(n), C.c_int(predict_type), C.c_int(start_iteration),
                                                C.c_int(num_iteration), C.byref(need)), "Booster CalcNumPredict")
        out = np.zeros(max(need.value, 1), dtype=np.float64)
        out_len = C.c_int64()
        validate(lib.LGBM_BoosterPredictForMat(self.booster_handle, x.ctypes.data_as(C.c_void_p), C.c_int(1), C.c_int32(n), C.c_int32(ncol),
                                               C.c_int(1), C.c_int(predict_type), C.c_int(start_iteration), C.c_int(num_iteration),
                                               ("max_bin=255 predict_disable_shape_check=%s" % ("true" if disable_shape_check else "false")).encode(),
                                               C.byref(out_len), out.ctypes.data_as(C.POINTER(C.c_double))), "Booster Predict")
        out = out[:out_len.value]
        return out.reshape(n, -1) if predict_type in (self.PREDICT_LEAF, self.PREDICT_CONTRIB) else out

    def score(self, row, raw=False, classification=None, disable_shape_check=False):
        """score (:394-405 -> predictForMat :520-557): one row through LGBM_BoosterPredictForMatSingleRow.  Without `classification`
        the single native value; with it the reference's array — one value per class, a binary model's single value expanded to
        [-raw, raw] or [1 - p, p] (predScoreToArray :559-575)."""
        lib = lightgbmlib()
        x = np.ascontiguousarray(row, dtype=np.float64).reshape(-1)
        k = self.num_classes()
        out, out_len = np.zeros(k), C.c_int64()
        validate(lib.LGBM_BoosterPredictForMatSingleRow(self.booster_handle, x.ctypes.data_as(C.c_void_p), C.c_int(1), C.c_int(len(x)), C.c_int(1),
                                                        C.c_int(self.PREDICT_RAW if raw else self.PREDICT_NORMAL), C.c_int(self.start_iteration),
                                                        C.c_int(self.num_iterations_predict),
                                                        ("max_bin=255 predict_disable_shape_check=%s" % ("true" if disable_shape_check else "false")).encode(),
                                                        C.byref(out_len), out.ctypes.data_as(C.POINTER(C.c_double))), "Booster Predict")
        if classification is None:
            return float(out[0])
        if classification and k == 1:
            return [-out[0], out[0]] if raw else [1 - out[0], out[0]]
        return out.tolist()

    def predict_leaf(self, row):
        """predictLeaf (:407-416): the leaf index of one row in every tree."""
        return self.predict_for_mat(row, self.PREDICT_LEAF, self.start_iteration, self.num_iterations_predict)[0].tolist()

    def features_shap(self, row):
        """featuresShap (:418-427): (numFeatures + 1) values per class — the TreeSHAP contributions and the expected value."""
        return self.predict_for_mat(row, self.PREDICT_CONTRIB, self.start_iteration, self.num_iterations_predict)[0].tolist()

    # prediction window (:433-451): first iteration used and how many (<= 0: all of them from the start iteration on)
    start_iteration, num_iterations_predict, best_iteration = 0, -1, -1

    def set_start_iteration(self, start_iteration):
        self.start_iteration = int(start_iteration)

    def set_num_iterations(self, num_iterations):
        self.num_iterations_predict = int(num_iterations)

    def set_best_iteration(self, best_iteration):
        self.best_iteration = int(best_iteration)
        self.num_iterations_predict = int(best_iteration)

    def get_native_model(self):
        """getNativeModel (:471-473)."""
        return self.model_str if self.model_str is not None else self.save_to_string()

    def save_native_model(self, filename, overwrite=True):
        """saveNativeModel (:458-468): the model text as a file; an existing file is an error unless overwrite."""
        if not filename:
            raise ValueError("filename should not be empty or null.")
        with open(filename, "w" if overwrite else "x") as f:
            f.write(self.get_native_model())

    def num_model_per_iteration(self):
        out = C.c_int()
        validate(lightgbmlib().LGBM_BoosterNumModelPerIteration(self.booster_handle, C.byref(out)), "Booster models per iteration")
        return out.value

    def num_total_iterations(self):
        return self.num_total_model() // self.num_model_per_iteration()      # :518

    def update_one_iteration(self):
        fin = C.c_int()
        validate(lightgbmlib().LGBM_BoosterUpdateOneIter(self.booster_handle, C.byref(fin)), "Booster Update One Iter")
        self.is_finished = fin.value == 1
        return self.is_finished

    def dump_model(self, importance_type="split"):
        """dumpModel (:480-499): the model as a JSON string (LGBM_BoosterDumpModel, grow-and-retry like saveToString)."""
        lib = lightgbmlib()
        n = C.c_int64()
        buf = C.create_string_buffer(1 << 20)
        args = (self.booster_handle, C.c_int(0), C.c_int(-1), C.c_int(0 if importance_type == "split" else 1))
        validate(lib.LGBM_BoosterDumpModel(*args, C.c_int64(len(buf)), C.byref(n), buf), "Booster Dump Model")
        if n.value > len(buf):
            buf = C.create_string_buffer(n.value)
            validate(lib.LGBM_BoosterDumpModel(*args, C.c_int64(len(buf)), C.byref(n), buf), "Booster Dump Model")
        return buf.value.decode()

    def merge_booster(self, model):
        """mergeBooster (:254-258): the trees of the serialized model go in front of this booster's (continued training)."""
        other, n_iter = C.c_void_p(), C.c_int()
        lib = lightgbmlib()
        validate(lib.LGBM_BoosterLoadModelFromString(model.encode(), C.byref(n_iter), C.byref(other)), "Booster LoadFromString")
        try:
            validate(lib.LGBM_BoosterMerge(self.booster_handle, other), "Booster Merge")
        finally:
            lib.LGBM_BoosterFree(other)

    def update_one_iteration_custom(self, gradient, hessian):
        """updateOneIterationCustom (:371-392): one iteration from caller-supplied gradients (float32, one per training row and
        class, class after class)."""
        g = np.ascontiguousarray(gradient, dtype=np.float32).reshape(-1)
        h = np.ascontiguousarray(hessian, dtype=np.float32).reshape(-1)
        want = self._train_rows * self.num_classes()
        if len(g) != want or len(h) != want:
            raise ValueError("gradient / hessian must have one value per training row and class (%d)" % want)
        fin = C.c_int()
        validate(lightgbmlib().LGBM_BoosterUpdateOneIterCustom(self.booster_handle, g.ctypes.data_as(C.POINTER(C.c_float)),
                                                               h.ctypes.data_as(C.POINTER(C.c_float)), C.byref(fin)),
                 "Booster Update One Iter Custom")
        self.is_finished = fin.value == 1
        return self.is_finished

    def reset_parameter(self, new_parameters):
        validate(lightgbmlib().LGBM_BoosterResetParameter(self.booster_handle, new_parameters.encode()), "Booster Reset learning_rate Param")

    def save_to_string(self, upper_iteration=None):
        """saveToString: LGBM_BoosterSaveModelToString with grow-and-retry (what the SWIG wrapper does)."""
        lib = lightgbmlib()
        n_iter = -1 if upper_iteration is None else int(upper_iteration)
        # first guess: ~4 KB per 31-leaf tree, so that the usual model is written once and not generated a second time for the retry
        size = max(1 << 20, self.num_total_model() * 4096 + (1 << 16))
        while True:
            buf = C.create_string_buffer(size)
            out_len = C.c_int64()
            validate(lib.LGBM_BoosterSaveModelToString(self.booster_handle, 0, n_iter, 0, C.c_int64(size), C.byref(out_len), buf),
                     "Booster save to string")
            if out_len.value <= size:
                return buf.value.decode()
            size = out_len.value

    def add_valid_data(self, dataset):
        """LGBM_BoosterAddValidData (BasePartitionTask.scala: the validation dataset is attached before training starts)."""
        validate(lightgbmlib().LGBM_BoosterAddValidData(self.booster_handle, dataset.dataset_ptr), "Add Validation Dataset")

    def get_eval_names(self):
        """getEvalNames (:283-297)."""
        lib = lightgbmlib()
        n = C.c_int()
        validate(lib.LGBM_BoosterGetEvalCounts(self.booster_handle, C.byref(n)), "Booster Get Eval Counts")
        bufs = [C.create_string_buffer(256) for _ in range(n.value)]
        arr = (C.c_char_p * max(n.value, 1))(*[C.cast(b, C.c_char_p) for b in bufs])
        out_n, need = C.c_int(), C.c_size_t()
        validate(lib.LGBM_BoosterGetEvalNames(self.booster_handle, C.c_int(n.value), C.byref(out_n), C.c_size_t(256), C.byref(need), arr),
                 "Booster Get Eval Names")
        return [b.value.decode() for b in bufs[:out_n.value]]

    def get_eval_results(self, eval_names, data_index):
        """getEvalResults (:300-314): [(name, value)] for data_index 0 (training data) or i (i-th validation set)."""
        out = np.zeros(max(len(eval_names), 1), dtype=np.float64)
        n = C.c_int()
        validate(lightgbmlib().LGBM_BoosterGetEval(self.booster_handle, C.c_int(data_index), C.byref(n), out.ctypes.data_as(C.POINTER(C.c_double))),
                 "Booster Get Eval Results for data index: %d" % data_index)
        return list(zip(eval_names, out[:len(eval_names)].tolist()))

    def num_classes(self):
        """getNumClasses (LightGBMBooster.scala:160-168): 1 unless the objective is multiclass."""
        k = C.c_int()
        validate(lightgbmlib().LGBM_BoosterGetNumClasses(self.booster_handle, C.byref(k)), "Booster NumClasses")
        return k.value

    def inner_predict(self, data_index=0):
        """Training scores; multiclass: class-major [num_class * num_data]."""
        lib = lightgbmlib()
        out_len = C.c_int64()
        out = np.zeros(self._train_rows * self.num_classes(), dtype=np.float64)
        validate(lib.LGBM_BoosterGetPredict(self.booster_handle, C.c_int(data_index), C.byref(out_len), out.ctypes.data_as(C.POINTER(C.c_double))),
                 "Booster get predict")
        return out[:out_len.value]

    def get_feature_importances(self, importance_type="split"):
        """getFeatureImportances (:501-508)."""
        out = np.zeros(self.num_features(), dtype=np.float64)
        validate(lightgbmlib().LGBM_BoosterFeatureImportance(self.booster_handle, C.c_int(-1), C.c_int(0 if importance_type == "split" else 1),
                                                             out.ctypes.data_as(C.POINTER(C.c_double))), "Booster FeatureImportance")
        return out.tolist()

    def num_total_model(self):
        out = C.c_int()
        validate(lightgbmlib().LGBM_BoosterNumberOfTotalModel(self.booster_handle, C.byref(out)), "Booster NumberOfTotalModel")
        return out.value

    def num_features(self):
        out = C.c_int()
        validate(lightgbmlib().LGBM_BoosterGetNumFeature(self.booster_handle, C.byref(out)), "Booster NumFeature")
        return out.value

    def stats(self):
        rows, ms, launches, hl = C.c_int64(), C.c_double(), C.c_int64(), C.c_int64()
        validate(lightgbmlib().LGBM_B200_BoosterGetStats(self.booster_handle, C.byref(rows), C.byref(ms), C.byref(launches), C.byref(hl)), "stats")
        r, e, sp, hk = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
        validate(lightgbmlib().LGBM_B200_BoosterGetGrowthStats(self.booster_handle, C.byref(r), C.byref(e), C.byref(sp), C.byref(hk)), "stats")
        return dict(hist_rows_visited=rows.value, hist_ms=ms.value, kernel_launches=launches.value, hist_launches=hl.value,
                    growth_rounds=r.value, growth_expansions=e.value, growth_splits=sp.value, hist_kernel_launches=hk.value)

    def timer_start(self):
        validate(lightgbmlib().LGBM_B200_BoosterTimerStart(self.booster_handle), "timer")

    def timer_stop(self):
        ms = C.c_float()
        validate(lightgbmlib().LGBM_B200_BoosterTimerStop(self.booster_handle, C.byref(ms)), "timer")
        return ms.value

    def set_profile(self, on):
        """Per-kernel-type CUDA-event timing of the training kernels (measurement only: it serialises the launches)."""
        validate(lightgbmlib().LGBM_B200_BoosterSetProfile(self.booster_handle, C.c_int(1 if on else 0)), "profile")

    def get_profile(self):
        """{"setup_ms": {...}, "kernels": {name: {"launches", "ms"}}, "partition_rows", "exchange"}"""
        import json
        n = C.c_int64()
        buf = C.create_string_buffer(1 << 14)
        validate(lightgbmlib().LGBM_B200_BoosterGetProfile(self.booster_handle, C.c_int64(len(buf)), C.byref(n), buf), "profile")
        return json.loads(buf.value.decode())

    def set_hist_timing(self, on):
        validate(lightgbmlib().LGBM_B200_BoosterSetHistTiming(self.booster_handle, C.c_int(1 if on else 0)), "hist timing")

    def free_native_memory(self):
        if self.booster_handle is not None:
            validate(lightgbmlib().LGBM_BoosterFree(self.booster_handle), "Finalize Booster")
            self.booster_handle = None
