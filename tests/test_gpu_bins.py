"""K1 (bin-boundary search) and K2 (row binning) through the C ABI vs the oracle and the committed golden fixture: bit-exact."""
import hashlib
import json
import os
import threading

import numpy as np
import pytest

from helpers import NativeDataset, edge_case_matrix, oracle_dataset
from synapseml_b200.data import synthetic as S

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def assert_same_mappers(nd, oref, nfeat):
    for j in range(nfeat):
        assert nd.feature_info(j) == oref.feature_info(j), j
        assert np.array_equal(nd.bounds(j), oref.bin_upper_bounds(j), equal_nan=True), j


def assert_same_bins(nd, ods):
    assert nd.num_used() == ods.num_used_feature()
    for j in range(nd.num_used()):
        assert np.array_equal(nd.bin_column(j).astype(np.uint16), ods.bin_column(j)), j
    rm = nd.row_major()
    cm = np.stack([nd.bin_column(j) for j in range(nd.num_used())], axis=1)
    assert np.array_equal(rm[:, :nd.num_used()], cm)
    assert not rm[:, nd.num_used():].any()


@pytest.mark.parametrize("max_bin", [255, 63, 15, 2])
def test_edge_case_columns(native, oracle, max_bin):
    X = edge_case_matrix(6000, 3)
    params = "max_bin=%d min_data_in_leaf=20" % max_bin
    nd = NativeDataset(X, np.zeros(len(X)), params, sample_rows=4000, micro_batches=[1, 7, 100, 1000, 333])
    oref, (ods,), _ = oracle_dataset(oracle, X, np.zeros(len(X), dtype=np.float32), params, sample_rows=4000)
    assert_same_mappers(nd, oref, X.shape[1])
    assert_same_bins(nd, ods)
    nd.free()


def test_golden_fixture(native):
    g = json.load(open(os.path.join(GOLDEN, "edge_bins.json")))
    X = edge_case_matrix(g["n"], g["seed"])
    nd = NativeDataset(X, np.zeros(len(X)), g["params"], sample_rows=g["sample"], micro_batches=[4096, 5])
    for j, feat in enumerate(g["features"]):
        info = nd.feature_info(j)
        for k in ("num_bin", "missing_type", "default_bin", "most_freq_bin", "is_trivial"):
            assert info[k] == feat[k], (j, k)
        assert ["nan" if np.isnan(x) else float.hex(float(x)) for x in nd.bounds(j)] == feat["bounds_hex"], j
    assert nd.num_used() == g["num_used"]
    cols = np.stack([nd.bin_column(j) for j in range(nd.num_used())])
    assert hashlib.sha256(cols.tobytes()).hexdigest() == g["bins_sha256"]
    assert cols.T[:64].astype(int).tolist() == g["first_rows"]
    nd.free()


@pytest.mark.parametrize("n,f", [(50000, 28), (20000, 50), (6000, 130)])
def test_shapes_cover_all_three_row_record_layouts(native, oracle, n, f):
    X, y = S.higgs_shape(n, f=f)
    params = "max_bin=255 min_data_in_leaf=20"
    nd = NativeDataset(X, y, params, micro_batches=[4096, 4097, 100])
    oref, (ods,), _ = oracle_dataset(oracle, X, y, params)
    assert_same_mappers(nd, oref, f)
    assert_same_bins(nd, ods)
    nd.free()


def test_filter_count_uses_distributed_row_count(native, oracle):
    """min_data_in_leaf * sample / num_dist_row decides which sparse features are dropped."""
    rng = np.random.Generator(np.random.PCG64(4))
    n = 5000
    X = np.stack([rng.standard_normal(n), np.where(rng.random(n) < 0.002, 1.0, 0.0)], axis=1)
    for num_dist in (n, 50 * n):
        params = "max_bin=255 min_data_in_leaf=20"
        nd = NativeDataset(X, np.zeros(n), params, num_dist_rows=num_dist)
        oref, _, _ = oracle_dataset(oracle, X, np.zeros(n, dtype=np.float32), params, num_dist_rows=num_dist)
        assert_same_mappers(nd, oref, 2)
        nd.free()


def test_concurrent_pushes_with_distinct_tid(native, oracle):
    """N task threads push disjoint row ranges into one dataset handle (BasePartitionTask.scala:84-112)."""
    X, y = S.higgs_shape(40000)
    params = "max_bin=255 min_data_in_leaf=20"
    from synapseml_b200.lightgbm.dataset import ReferenceDatasetUtils
    from synapseml_b200.lightgbm.native import validate
    ser = ReferenceDatasetUtils.create_reference_dataset_from_sample(params, "features", len(y), X.shape[1], X)
    nd = NativeDataset.__new__(NativeDataset)
    nd.lib, nd.n, nd.f = native, len(y), X.shape[1]
    nd.handle = ReferenceDatasetUtils.get_initialized_reference_dataset(ser, len(y), 1, params, False, False, False, 4)
    bounds = [0, 9000, 9001, 25000, 40000]
    errs = []

    def work(tid):
        try:
            for s in range(bounds[tid], bounds[tid + 1], 777):
                e = min(bounds[tid + 1], s + 777)
                nd.push(X[s:e], s, y[s:e], None, None, None, tid)
        except Exception as ex:  # noqa: BLE001
            errs.append(ex)
    ts = [threading.Thread(target=work, args=(t,)) for t in range(4)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errs, errs
    validate(native.LGBM_DatasetMarkFinished(nd.handle))
    _, (ods,), _ = oracle_dataset(oracle, X, y, params)
    assert_same_bins(nd, ods)
    nd.free()


def test_serialized_reference_round_trip_and_errors(native):
    import ctypes as C
    from synapseml_b200.lightgbm.native import LightGBMNativeError, validate
    X, y = S.higgs_shape(3000)
    nd = NativeDataset(X, y, "max_bin=31")
    nd2 = NativeDataset(X, y, "max_bin=31", serialized=nd.serialized)
    for j in range(X.shape[1]):
        assert nd.feature_info(j) == nd2.feature_info(j) and np.array_equal(nd.bounds(j), nd2.bounds(j))
    # per-byte accessor of the reference (ReferenceDatasetUtils.scala:99-119) agrees with the bulk copy
    bb, blen = C.c_void_p(), C.c_int32()
    validate(native.LGBM_DatasetSerializeReferenceToBinary(nd.handle, C.byref(bb), C.byref(blen)))
    v = C.c_uint8()
    for i in (0, 1, blen.value // 2, blen.value - 1):
        validate(native.LGBM_ByteBufferGetAt(bb, i, C.byref(v)))
        assert v.value == nd.serialized[i]
    assert native.LGBM_ByteBufferGetAt(bb, blen.value, C.byref(v)) == -1
    validate(native.LGBM_ByteBufferFree(bb))
    out = C.c_void_p()
    assert native.LGBM_DatasetCreateFromSerializedReference(b"garbage-bytes", 13, C.c_int64(10), 1, b"", C.byref(out)) == -1
    # wrong column count / out-of-range rows are rejected with a message, not a crash
    with pytest.raises(LightGBMNativeError):
        bad = np.zeros((4, X.shape[1] + 1))
        validate(native.LGBM_DatasetPushRowsWithMetadata(nd2.handle, bad.ctypes.data_as(C.c_void_p), 1, 4, bad.shape[1], 0, None, None, None, None, 0))
    nd.free()
    nd2.free()


def test_max_bin_by_feature(native, oracle):
    """max_bin_by_feature (GeneralParams.maxBinByFeature, params/TrainParams.scala) at dataset construction: one max_bin per column, some
    above 255 (several uint8 columns, csrc/dataset.h), some tiny."""
    X, y = S.higgs_shape(30000, f=8)
    params = "max_bin=255 min_data_in_leaf=20 max_bin_by_feature=4,300,17,3,255,1000,64,2"
    nd = NativeDataset(X, y, params, micro_batches=[4096, 100])
    oref, (ods,), _ = oracle_dataset(oracle, X, y, params)
    assert_same_mappers(nd, oref, 8)
    assert_same_bins_logical(nd, ods)
    bins = [nd.feature_info(j)["num_bin"] for j in range(8)]
    assert bins[0] <= 4 and bins[1] > 255 and bins[5] > 255 and bins[2] <= 17
    nd.free()


def assert_same_bins_logical(nd, ods):
    assert nd.num_used() == ods.num_used_feature()
    for j in range(nd.num_used()):
        assert np.array_equal(nd.bin_column(j).astype(np.uint16), ods.bin_column(j)), j


@pytest.mark.parametrize("params", ["max_bin=255 min_data_in_leaf=20", "max_bin=15 min_data_in_leaf=5 zero_as_missing=true",
                                    "max_bin=600 min_data_in_leaf=20 categorical_feature=7,10", "max_bin=63 use_missing=false"])
def test_row_major_sample_gives_the_reference_of_the_column_arrays(native, params):
    # LGBM_B200_DatasetCreateFromSampledRows (zeros dropped and columns gathered on the device) against
    # LGBM_DatasetCreateFromSampledColumn fed with SampledData's host-built arrays: the serialized bin mappers must be the same bytes
    from synapseml_b200.lightgbm.dataset import ReferenceDatasetUtils
    for X in (edge_case_matrix(6000, 3), S.higgs_shape(20000)[0], np.zeros((50, 3)), -np.zeros((40, 2)),
              np.random.Generator(np.random.PCG64(1)).normal(size=(3000, 70))):
        if "categorical" in params and X.shape[1] < 11:
            continue
        out = {}
        for flag in (True, False):
            ReferenceDatasetUtils.row_major_sample = flag
            try:
                out[flag] = ReferenceDatasetUtils.create_reference_dataset_from_sample(params, "features", 10 * len(X), X.shape[1], X)
            finally:
                ReferenceDatasetUtils.row_major_sample = True
        assert out[True] == out[False] and len(out[True]) > 16, (params, X.shape)
