"""Where does dataset creation time go?  (GPU box only.)"""
import sys, time, threading
import numpy as np
sys.path.insert(0, ".")
from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LocalDataFrame
from synapseml_b200.lightgbm.dataset import ReferenceDatasetUtils
from synapseml_b200.lightgbm.partition_task import StreamingPartitionTask
from synapseml_b200.lightgbm.native import lightgbmlib, validate
from synapseml_b200.lightgbm.params import dataset_creation_params

N = 10_000_000
X, y = S.higgs_shape(N)
y = y.astype(np.float32)
ds_params = dataset_creation_params(255, 200000, 20, 0, data_random_seed=0)
t0 = time.time()
ser = ReferenceDatasetUtils.create_reference_dataset_from_sample(ds_params, "features", N, 28, X[:200000])
print("reference dataset from sample (1st, includes CUDA init): %.3f s" % (time.time() - t0))
t0 = time.time()
ser = ReferenceDatasetUtils.create_reference_dataset_from_sample(ds_params, "features", N, 28, X[:200000])
print("reference dataset from sample (2nd): %.3f s" % (time.time() - t0))
for nt in (1, 4, 8, 16):
    for rep in range(2):
        t0 = time.time()
        h = ReferenceDatasetUtils.get_initialized_reference_dataset(ser, N, 1, ds_params, False, False, False, nt)
        t1 = time.time()
        df = LocalDataFrame.from_arrays(X, y, num_partitions=nt)
        offs = np.concatenate([[0], np.cumsum([len(p["label"]) for p in df.partitions])])
        ths = [threading.Thread(target=StreamingPartitionTask.insert_rows_into_training_dataset,
                                args=(h, p, int(offs[i]), 1 << 16, i, None, False, False, False)) for i, p in enumerate(df.partitions)]
        [t.start() for t in ths]; [t.join() for t in ths]
        t2 = time.time()
        validate(lightgbmlib().LGBM_DatasetMarkFinished(h), "finish")
        t3 = time.time()
        validate(lightgbmlib().LGBM_DatasetFree(h), "free")
        t4 = time.time()
        print("threads %2d: init %.3f  push %.3f (%.1f GB/s)  finish %.3f  free %.3f" % (nt, t1 - t0, t2 - t1, X.nbytes / (t2 - t1) / 1e9, t3 - t2, t4 - t3))
