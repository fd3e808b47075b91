"""Where does the bin-boundary phase go?  (GPU box only.)"""
import sys, time, ctypes as C
import numpy as np
sys.path.insert(0, ".")
from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.dataset import SampledData
from synapseml_b200.lightgbm.native import lightgbmlib, validate
from synapseml_b200.lightgbm.params import dataset_creation_params

X, y = S.higgs_shape(200000)
ds_params = dataset_creation_params(255, 200000, 20, 0, data_random_seed=0)
lib = lightgbmlib()
for rep in range(3):
    t0 = time.time()
    sd = SampledData(len(X), 28); sd.push_rows(X)
    t1 = time.time()
    h = C.c_void_p()
    validate(lib.LGBM_DatasetCreateFromSampledColumn(sd.get_sample_data(), sd.get_sample_indices(), C.c_int32(28), sd.get_row_counts(),
                                                     C.c_int32(len(X)), C.c_int32(1), C.c_int64(10_000_000), ds_params.encode(), C.byref(h)), "create")
    t2 = time.time()
    validate(lib.LGBM_DatasetFree(h), "free")
    print("SampledData %.3f s   CreateFromSampledColumn %.3f s" % (t1 - t0, t2 - t1))
