"""Scratch GPU check used during bring-up (run by hand: python experiments/quick_gpu.py)."""
import ctypes as C, sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as O
from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.native import lightgbmlib, validate

lib = lightgbmlib()
N = int(os.environ.get("N", 200000)); ITERS = int(os.environ.get("ITERS", 10))
X, y = S.higgs_shape(N)
params = os.environ.get("PARAMS", "objective=binary num_leaves=31 learning_rate=0.1 max_bin=255 min_data_in_leaf=20 use_quantized_grad=true stochastic_rounding=false verbosity=1")
sample = X[:min(200000, N)]
cols, idxs, counts = O.sampled_columns(sample)
f = len(cols)
dpp = (C.POINTER(C.c_double) * f)(*[c.ctypes.data_as(C.POINTER(C.c_double)) for c in cols])
ipp = (C.POINTER(C.c_int) * f)(*[i.ctypes.data_as(C.POINTER(C.c_int)) for i in idxs])
ref = C.c_void_p()
t = time.time()
validate(lib.LGBM_DatasetCreateFromSampledColumn(dpp, ipp, C.c_int32(f), counts.ctypes.data_as(C.POINTER(C.c_int)), C.c_int32(sample.shape[0]), C.c_int32(1), C.c_int64(N), params.encode(), C.byref(ref)))
print("K1 time", time.time() - t)
bb = C.c_void_p(); blen = C.c_int32()
validate(lib.LGBM_DatasetSerializeReferenceToBinary(ref, C.byref(bb), C.byref(blen)))
buf = (C.c_uint8 * blen.value)()
validate(lib.LGBM_ByteBufferCopy(bb, buf, blen))
validate(lib.LGBM_ByteBufferFree(bb))
ds = C.c_void_p()
validate(lib.LGBM_DatasetCreateFromSerializedReference(buf, blen, C.c_int64(N), C.c_int32(1), params.encode(), C.byref(ds)))
validate(lib.LGBM_DatasetInitStreaming(ds, 0, 0, 0, 1, 1, 1))
t = time.time()
mb = 50000
for s in range(0, N, mb):
    xb = np.ascontiguousarray(X[s:s + mb]); yb = np.ascontiguousarray(y[s:s + mb])
    validate(lib.LGBM_DatasetPushRowsWithMetadata(ds, xb.ctypes.data_as(C.c_void_p), 1, C.c_int32(xb.shape[0]), C.c_int32(f), C.c_int32(s), yb.ctypes.data_as(C.POINTER(C.c_float)), None, None, None, 0))
validate(lib.LGBM_DatasetMarkFinished(ds))
print("push time", time.time() - t)
# oracle
oref = O.Dataset.from_sample(sample, 1, N, params)
ods = oref.create_by_reference(N); ods.init_streaming(); ods.push_rows(X, 0, y); ods.mark_finished()
# bounds parity
nb_bad = 0
for j in range(f):
    info = oref.feature_info(j)
    nb, mt, db, mf, tr = (C.c_int() for _ in range(5)); mn, mx = C.c_double(), C.c_double()
    validate(lib.LGBM_B200_DatasetGetFeatureInfo(ds, j, C.byref(nb), C.byref(mt), C.byref(db), C.byref(mf), C.byref(tr), C.byref(mn), C.byref(mx)))
    got = dict(num_bin=nb.value, missing_type=mt.value, default_bin=db.value, most_freq_bin=mf.value, is_trivial=bool(tr.value), min_val=mn.value, max_val=mx.value)
    if got != info: print("feature", j, "info mismatch", got, info); nb_bad += 1
    ub = np.zeros(nb.value); validate(lib.LGBM_B200_DatasetGetBinUpperBounds(ds, j, ub.ctypes.data_as(C.POINTER(C.c_double))))
    if not np.array_equal(ub, oref.bin_upper_bounds(j), equal_nan=True): print("feature", j, "bounds mismatch"); nb_bad += 1
print("bin mapper mismatches:", nb_bad)
nu = C.c_int(); validate(lib.LGBM_B200_DatasetGetNumUsedFeature(ds, C.byref(nu)))
bad = 0
for j in range(nu.value):
    col = np.zeros(N, dtype=np.uint8); validate(lib.LGBM_B200_DatasetCopyBinColumn(ds, j, col.ctypes.data_as(C.POINTER(C.c_uint8))))
    oc = ods.bin_column(j)
    if not np.array_equal(col.astype(np.uint16), oc): bad += 1; print("bin column", j, "mismatch", np.flatnonzero(col != oc)[:5])
print("bin column mismatches:", bad)
stride = C.c_int(); validate(lib.LGBM_B200_DatasetCopyRowMajor(ds, 0, 0, C.byref(stride), None))
rmm = np.zeros((min(N, 5000), stride.value), dtype=np.uint8); validate(lib.LGBM_B200_DatasetCopyRowMajor(ds, 0, rmm.shape[0], C.byref(stride), rmm.ctypes.data_as(C.POINTER(C.c_uint8))))
cm0 = np.stack([ods.bin_column(j)[:rmm.shape[0]] for j in range(nu.value)], axis=1)
print("row-major ok:", np.array_equal(rmm[:, :nu.value].astype(np.uint16), cm0))
# train
bst = C.c_void_p(); validate(lib.LGBM_BoosterCreate(ds, params.encode(), C.byref(bst)))
obst = O.Booster([ods], params)
fin = C.c_int()
t = time.time()
for it in range(ITERS):
    validate(lib.LGBM_BoosterUpdateOneIter(bst, C.byref(fin)))
t_gpu = time.time() - t
t = time.time()
for it in range(ITERS): obst.update_one_iter()
t_cpu = time.time() - t
print("gpu iters %.3fs  oracle iters %.3fs" % (t_gpu, t_cpu))
def model(b):
    n = C.c_int64(); sb = C.create_string_buffer(1 << 20)
    validate(lib.LGBM_BoosterSaveModelToString(b, 0, -1, 0, C.c_int64(len(sb)), C.byref(n), sb))
    return sb.value.decode()
m1, m2 = model(bst), obst.save_model_to_string()
def strip_params(m): return m.split("parameters:")[0]
print("model equal:", strip_params(m1) == strip_params(m2))
if strip_params(m1) != strip_params(m2):
    a, b = strip_params(m1).split("\n"), strip_params(m2).split("\n")
    for i, (u, v) in enumerate(zip(a, b)):
        if u != v: print("first diff line", i, "\n GPU:", u[:300], "\n ORC:", v[:300]); break
sc = np.zeros(N); n64 = C.c_int64(); validate(lib.LGBM_BoosterGetPredict(bst, 0, C.byref(n64), sc.ctypes.data_as(C.POINTER(C.c_double))))
osc = obst.scores()
print("score max abs diff", np.abs(sc - osc).max(), "rel", np.abs(sc - osc).max() / np.abs(osc).max())
