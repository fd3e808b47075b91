"""LightGBMBooster: mirror of lightgbm/src/main/scala/com/microsoft/azure/synapse/ml/lightgbm/booster/LightGBMBooster.scala
(training subset: create :241, updateOneIteration :355-365, resetParameter :316-323, saveToString :272-278,
getPredict :331-350, free :155) and TrainUtils.createBooster (TrainUtils.scala:16-29)."""
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
            self._train_rows = training_dataset.num_data()

    def update_one_iteration(self):
        fin = C.c_int()
        validate(lightgbmlib().LGBM_BoosterUpdateOneIter(self.booster_handle, C.byref(fin)), "Booster Update One Iter")
        self.is_finished = fin.value == 1
        return self.is_finished

    def reset_parameter(self, new_parameters):
        validate(lightgbmlib().LGBM_BoosterResetParameter(self.booster_handle, new_parameters.encode()), "Booster Reset learning_rate Param")

    def save_to_string(self, upper_iteration=None):
        """saveToString: LGBM_BoosterSaveModelToString with grow-and-retry (what the SWIG wrapper does)."""
        lib = lightgbmlib()
        n_iter = -1 if upper_iteration is None else int(upper_iteration)
        size = 1 << 20
        while True:
            buf = C.create_string_buffer(size)
            out_len = C.c_int64()
            validate(lib.LGBM_BoosterSaveModelToString(self.booster_handle, 0, n_iter, 0, C.c_int64(size), C.byref(out_len), buf),
                     "Booster save to string")
            if out_len.value <= size:
                return buf.value.decode()
            size = out_len.value

    def inner_predict(self, data_index=0):
        lib = lightgbmlib()
        out_len = C.c_int64()
        out = np.zeros(self._train_rows, dtype=np.float64)
        validate(lib.LGBM_BoosterGetPredict(self.booster_handle, C.c_int(data_index), C.byref(out_len), out.ctypes.data_as(C.POINTER(C.c_double))),
                 "Booster get predict")
        return out[:out_len.value]

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

    def set_hist_timing(self, on):
        validate(lightgbmlib().LGBM_B200_BoosterSetHistTiming(self.booster_handle, C.c_int(1 if on else 0)), "hist timing")

    def free_native_memory(self):
        if self.booster_handle is not None:
            validate(lightgbmlib().LGBM_BoosterFree(self.booster_handle), "Finalize Booster")
            self.booster_handle = None
