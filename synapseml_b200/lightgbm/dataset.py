"""Host-side dataset layer above the C ABI.

Mirrors (same names, argument meaning and error behaviour):
  SampledData                  lightgbm/.../dataset/SampledData.scala:26-96
  ReferenceDatasetUtils        lightgbm/.../dataset/ReferenceDatasetUtils.scala:13-139
  LightGBMDataset              lightgbm/.../dataset/LightGBMDataset.scala:17-193
  DatasetUtils.countCardinality lightgbm/.../dataset/DatasetUtils.scala:15-37
  GroupIdManager               lightgbm/.../GroupIdManager.scala:15-45
"""
import ctypes as C

import numpy as np

from .native import C_API_DTYPE_FLOAT32, C_API_DTYPE_FLOAT64, C_API_DTYPE_INT32, lightgbmlib, validate


def count_cardinality(values):
    """DatasetUtils.countCardinality: run lengths of consecutive equal values (rows are sorted by group within a partition)."""
    out = []
    cur, cnt = None, 0
    for v in values:
        if cnt == 0:
            cur, cnt = v, 1
        elif v == cur:
            cnt += 1
        else:
            out.append(cnt)
            cur, cnt = v, 1
    out.append(cnt)
    return out


class GroupIdManager:
    """Maps group column values (int, long or string) to dense int ids in order of first appearance."""

    def __init__(self):
        self._ids = {}

    def get_unique_index_for_id(self, value):
        idx = self._ids.get(value)
        if idx is None:
            idx = len(self._ids)
            self._ids[value] = idx
        return idx


class SampledData:
    """Column-major, non-zero-only copy of the sampled rows: the input layout of LGBM_DatasetCreateFromSampledColumn."""

    def __init__(self, num_rows, num_cols):
        self.num_rows, self.num_cols = num_rows, num_cols
        # one [num_cols][num_rows] block per array: column j's buffer is row j (each pre-sized to num_rows, SampledData.scala:35-40)
        self._data = np.zeros((num_cols, num_rows), dtype=np.float64)
        self._idx = np.zeros((num_cols, num_rows), dtype=np.int32)
        self.sample_data = [self._data[j] for j in range(num_cols)]
        self.sample_indexes = [self._idx[j] for j in range(num_cols)]
        self.row_counts = np.zeros(num_cols, dtype=np.int32)
        self._next_row = 0

    def push_row(self, row, index):
        """SampledData.pushRow for a dense row (SampledData.scala:43-77): zeros are skipped, NaN is kept."""
        nz = np.flatnonzero(row != 0.0)
        for col in nz:
            k = self.row_counts[col]
            self.sample_data[col][k] = row[col]
            self.sample_indexes[col][k] = index
            self.row_counts[col] = k + 1

    def push_rows(self, rows, block=None):
        """Equivalent of calling push_row for rows 0..n-1 in order, done in cache-sized row blocks: a block is transposed once, then a
        column without zeros is one contiguous copy and a column with zeros a gather of its non-zero positions."""
        rows = np.asarray(rows, dtype=np.float64)
        n = rows.shape[0]
        base = self._next_row
        if block is None:      # ~2 MB of doubles per block (the transposed block stays in cache while its columns are copied), but
            block = int(min(65536, max(4096, (1 << 18) // max(self.num_cols, 1))))      # never so few rows that the per-column Python loop dominates on wide data
        for b in range(0, n, block):
            blk = rows[b:b + block]
            nb = blk.shape[0]
            k0 = int(self.row_counts[0]) if self.num_cols else 0
            if (self.row_counts == k0).all() and (blk != 0.0).all():      # dense data: all columns with one transposed copy
                self._data[:, k0:k0 + nb] = blk.T
                self._idx[:, k0:k0 + nb] = np.arange(base + b, base + b + nb, dtype=np.int32)
                self.row_counts += nb
                continue
            blk_t = np.ascontiguousarray(blk.T)
            full = (blk_t != 0.0).all(axis=1)
            index = None
            for col in range(self.num_cols):
                k = int(self.row_counts[col])
                if full[col]:
                    if index is None:
                        index = np.arange(base + b, base + b + nb, dtype=np.int32)
                    self._data[col, k:k + nb] = blk_t[col]
                    self._idx[col, k:k + nb] = index
                    self.row_counts[col] = k + nb
                    continue
                nz = np.flatnonzero(blk_t[col])      # NaN counts as non-zero, -0.0 as zero: as `!= 0.0` does
                if len(nz) == 0:
                    continue
                self._data[col, k:k + len(nz)] = blk_t[col, nz]
                self._idx[col, k:k + len(nz)] = nz + (base + b)
                self.row_counts[col] = k + len(nz)
        self._next_row = base + n

    def get_sample_data(self):
        return (C.POINTER(C.c_double) * self.num_cols)(*[a.ctypes.data_as(C.POINTER(C.c_double)) for a in self.sample_data])

    def get_sample_indices(self):
        return (C.POINTER(C.c_int) * self.num_cols)(*[a.ctypes.data_as(C.POINTER(C.c_int)) for a in self.sample_indexes])

    def get_row_counts(self):
        return self.row_counts.ctypes.data_as(C.POINTER(C.c_int))


class LightGBMDataset:
    """Handle wrapper (LightGBMDataset.scala)."""

    def __init__(self, handle):
        self.dataset_ptr = handle

    def _get_int(self, fn, what):
        out = C.c_int()
        validate(fn(self.dataset_ptr, C.byref(out)), what)
        return out.value

    def num_data(self):
        return self._get_int(lightgbmlib().LGBM_DatasetGetNumData, "DatasetGetNumData")

    def num_feature(self):
        return self._get_int(lightgbmlib().LGBM_DatasetGetNumFeature, "DatasetGetNumFeature")

    def validate_dataset(self):
        """LightGBMDataset.validateDataset (:71-83)."""
        if self.num_data() == 0:
            raise ValueError("No data in the dataset")
        if self.num_feature() == 0:
            raise ValueError("No features in the dataset")

    def _set_field(self, name, arr, dtype_code):
        validate(lightgbmlib().LGBM_DatasetSetField(self.dataset_ptr, name.encode(), arr.ctypes.data_as(C.c_void_p),
                                                    C.c_int(len(arr)), C.c_int(dtype_code)), "DatasetSetField")

    def add_float_field(self, field, name):
        self._set_field(name, np.ascontiguousarray(field, dtype=np.float32), C_API_DTYPE_FLOAT32)

    def add_double_field(self, field, name):
        self._set_field(name, np.ascontiguousarray(field, dtype=np.float64), C_API_DTYPE_FLOAT64)

    def add_int_field(self, field, name):
        self._set_field(name, np.ascontiguousarray(field, dtype=np.int32), C_API_DTYPE_INT32)

    def add_group_column(self, values):
        """LightGBMDataset.addGroupColumn (:171-176): cardinality-encode the (sorted) group column."""
        self.add_int_field(count_cardinality(list(values)), "group")

    def get_label(self):
        n, ptr, tp = C.c_int(), C.c_void_p(), C.c_int()
        validate(lightgbmlib().LGBM_DatasetGetField(self.dataset_ptr, b"label", C.byref(n), C.byref(ptr), C.byref(tp)), "DatasetGetField")
        return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_float)), shape=(n.value,)).copy()

    def set_feature_names(self, names, num_cols):
        """LightGBMDataset.setFeatureNames (:178-187)."""
        if names is None or len(names) == 0:
            return
        arr = (C.c_char_p * num_cols)(*[n.encode() for n in names])
        validate(lightgbmlib().LGBM_DatasetSetFeatureNames(self.dataset_ptr, arr, C.c_int(num_cols)), "Dataset set feature names")

    def close(self):
        if self.dataset_ptr is not None:
            validate(lightgbmlib().LGBM_DatasetFree(self.dataset_ptr), "Finalize Dataset")
            self.dataset_ptr = None


class ReferenceDatasetUtils:
    row_major_sample = True      # False: build SampledData on the host and call LGBM_DatasetCreateFromSampledColumn, as the reference does

    @staticmethod
    def create_reference_dataset_from_sample(dataset_params, feature_data_name, num_rows, num_cols, sampled_rows):
        """Driver side (ReferenceDatasetUtils.createReferenceDatasetFromSample, :14-71): sample -> bin mappers -> bytes."""
        del feature_data_name
        sampled_rows = np.ascontiguousarray(sampled_rows, dtype=np.float64)
        lib = lightgbmlib()
        handle = C.c_void_p()
        # num_local_row = 1 and num_dist_row = N, as the reference passes them (:44-45)
        if ReferenceDatasetUtils.row_major_sample and len(sampled_rows) > 0:
            # the collected rows go to the library as they are; it drops the zeros and gathers the columns on the device
            validate(lib.LGBM_B200_DatasetCreateFromSampledRows(sampled_rows.ctypes.data_as(C.POINTER(C.c_double)), C.c_int32(len(sampled_rows)),
                                                                C.c_int32(num_cols), C.c_int32(1), C.c_int64(num_rows),
                                                                dataset_params.encode(), C.byref(handle)), "Dataset create from samples")
        else:
            # the reference's own route: SampledData's column-major non-zero arrays (SampledData.scala:26-96)
            sampled = SampledData(len(sampled_rows), num_cols)
            sampled.push_rows(sampled_rows)
            validate(lib.LGBM_DatasetCreateFromSampledColumn(sampled.get_sample_data(), sampled.get_sample_indices(), C.c_int32(num_cols),
                                                             sampled.get_row_counts(), C.c_int32(len(sampled_rows)), C.c_int32(1),
                                                             C.c_int64(num_rows), dataset_params.encode(), C.byref(handle)),
                     "Dataset create from samples")
        try:
            buf, blen = C.c_void_p(), C.c_int32()
            validate(lib.LGBM_DatasetSerializeReferenceToBinary(handle, C.byref(buf), C.byref(blen)), "Serialize ref")
            try:
                out = (C.c_uint8 * blen.value)()
                # one bulk copy instead of the reference's per-byte LGBM_ByteBufferGetAt loop (:99-119)
                validate(lib.LGBM_ByteBufferCopy(buf, out, blen), "Buffer copy")
                return bytes(out)
            finally:
                validate(lib.LGBM_ByteBufferFree(buf), "Buffer free")
        finally:
            validate(lib.LGBM_DatasetFree(handle), "Dataset free")

    @staticmethod
    def get_initialized_reference_dataset(serialized_reference, num_rows, num_classes, dataset_params, has_weights,
                                          has_init_scores, has_groups, num_threads, max_omp_threads=16):
        """Executor side (getInitializedReferenceDataset, :73-97): empty num_rows-row dataset + streaming buffers."""
        lib = lightgbmlib()
        handle = C.c_void_p()
        buf = (C.c_uint8 * len(serialized_reference)).from_buffer_copy(serialized_reference)
        validate(lib.LGBM_DatasetCreateFromSerializedReference(buf, C.c_int32(len(serialized_reference)), C.c_int64(num_rows),
                                                               C.c_int32(num_classes), dataset_params.encode(), C.byref(handle)),
                 "Dataset create from reference")
        validate(lib.LGBM_DatasetInitStreaming(handle, C.c_int32(has_weights), C.c_int32(has_init_scores), C.c_int32(has_groups),
                                               C.c_int32(num_classes), C.c_int32(num_threads), C.c_int32(max_omp_threads)),
                 "LGBM_DatasetInitStreaming")
        return handle
