/*
 * C ABI of the B200-native LightGBM training library (lib_lightgbm.so built from synapseml_b200/csrc).
 *
 * Every entry point below has the name, argument order and meaning of the LightGBM C API function that the reference
 * binds through SWIG/JNI (com.microsoft.ml.lightgbm.lightgbmlib); the reference call site each one serves is cited
 * (paths relative to /root/reference/lightgbm/src/main/scala/com/microsoft/azure/synapse/ml/lightgbm/).
 * Conventions (reference: LightGBMUtils.scala:15-27): return 0 on success and -1 on error; the message is then available
 * from LGBM_GetLastError() on the calling thread.  Handles are opaque.  The caller owns every input buffer and may free
 * it as soon as the call returns (the library copies what it keeps).  No exception crosses this boundary.
 *
 * All compute runs on one B200 per process (device = parameter gpu_device_id, else $LOCAL_RANK, else 0); there is no CPU
 * path: a call that needs the GPU fails with an error when none is usable.
 */
#ifndef LIGHTGBM_B200_C_API_H_
#define LIGHTGBM_B200_C_API_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* DatasetHandle;
typedef void* BoosterHandle;
typedef void* ByteBufferHandle;

#define C_API_DTYPE_FLOAT32 (0)
#define C_API_DTYPE_FLOAT64 (1)
#define C_API_DTYPE_INT32 (2)
#define C_API_DTYPE_INT64 (3)

#define C_API_PREDICT_NORMAL (0)
#define C_API_PREDICT_RAW_SCORE (1)
#define C_API_PREDICT_LEAF_INDEX (2)
#define C_API_PREDICT_CONTRIB (3)

/* LightGBMUtils.scala:18,25 — thread-local message of the last failed call */
const char* LGBM_GetLastError(void);

/* ---------------------------------------------------------------- Dataset */

/* dataset/ReferenceDatasetUtils.scala:38-47 — bin-boundary search (kernel K1) on the driver-side sample laid out by
 * dataset/SampledData.scala:26-96 (column-major, non-zero values only). */
int LGBM_DatasetCreateFromSampledColumn(double** sample_data, int** sample_indices, int32_t ncol, const int* num_per_col,
                                        int32_t num_sample_row, int32_t num_local_row, int64_t num_dist_row,
                                        const char* parameters, DatasetHandle* out);

/* The same search from the sampled rows as they are — a dense row-major matrix [num_sample_row][ncol] — for callers that hold
 * the sample that way (LightGBMBase.getSampledRows, LightGBMBase.scala:724-749, collects rows): the zeros are dropped and the
 * columns gathered on the device, so the host never builds SampledData's per-column arrays.  Same result as the call above. */
int LGBM_B200_DatasetCreateFromSampledRows(const double* rows, int32_t num_sample_row, int32_t ncol, int32_t num_local_row,
                                           int64_t num_dist_row, const char* parameters, DatasetHandle* out);

/* dataset/ReferenceDatasetUtils.scala:52-55 — bin mappers as an opaque byte buffer (shipped driver -> executors) */
int LGBM_DatasetSerializeReferenceToBinary(DatasetHandle handle, ByteBufferHandle* out, int32_t* out_len);
/* dataset/ReferenceDatasetUtils.scala:107 / :115 */
int LGBM_ByteBufferGetAt(ByteBufferHandle handle, int32_t index, uint8_t* out_val);
int LGBM_ByteBufferFree(ByteBufferHandle handle);
/* bulk alternative to the per-byte LGBM_ByteBufferGetAt loop: copies the whole buffer */
int LGBM_ByteBufferCopy(ByteBufferHandle handle, uint8_t* out, int32_t out_capacity);

/* dataset/ReferenceDatasetUtils.scala:121-138 — empty num_row-row dataset with the serialized bin mappers */
int LGBM_DatasetCreateFromSerializedReference(const void* ref_buffer, int32_t ref_buffer_size, int64_t num_row,
                                              int32_t num_classes, const char* parameters, DatasetHandle* out);

/* StreamingPartitionTask.scala:349-364 (validation data) — empty dataset sharing another dataset's bin mappers */
int LGBM_DatasetCreateByReference(const DatasetHandle reference, int64_t num_total_row, DatasetHandle* out);

/* dataset/ReferenceDatasetUtils.scala:87-94 */
int LGBM_DatasetInitStreaming(DatasetHandle dataset, int32_t has_weights, int32_t has_init_scores, int32_t has_queries,
                              int32_t nclasses, int32_t nthreads, int32_t omp_max_threads);

/* StreamingPartitionTask.scala:220-231 — row binning (kernel K2).  Thread-safe for distinct tid and disjoint row ranges. */
int LGBM_DatasetPushRowsWithMetadata(DatasetHandle dataset, const void* data, int data_type, int32_t nrow, int32_t ncol,
                                     int32_t start_row, const float* label, const float* weight, const double* init_score,
                                     const int32_t* query, int32_t tid);

/* StreamingPartitionTask.scala:258-272 — sparse micro-batches (indptr_type C_API_DTYPE_INT32 = 2 / INT64 = 3; the reference passes
 * INT64 and float64 values).  Absent entries are 0.0.  Same threading rules as the dense push. */
int LGBM_DatasetPushRowsByCSRWithMetadata(DatasetHandle dataset, const void* indptr, int indptr_type, const int32_t* indices,
                                          const void* data, int data_type, int64_t nindptr, int64_t nelem, int64_t start_row,
                                          const float* label, const float* weight, const double* init_score, const int32_t* query,
                                          int32_t tid);

/* StreamingPartitionTask.scala:358 / :143,168 */
int LGBM_DatasetSetWaitForManualFinish(DatasetHandle dataset, int wait);
int LGBM_DatasetMarkFinished(DatasetHandle dataset);

/* dataset/DatasetAggregator.scala:413-421 (bulk mode): bin a whole row-major matrix; reference may be NULL */
int LGBM_DatasetCreateFromMat(const void* data, int data_type, int32_t nrow, int32_t ncol, int is_row_major,
                              const char* parameters, const DatasetHandle reference, DatasetHandle* out);
/* Bulk mode with sparse rows (LGBM/dataset/DatasetAggregator.scala: SparseAggregatedColumns.generateDataset): CSR in, same flow as
 * CreateFromMat (bin mappers from `reference` or from the first bin_construct_sample_cnt rows). */
int LGBM_DatasetCreateFromCSR(const void* indptr, int indptr_type, const int32_t* indices, const void* data, int data_type,
                              int64_t nindptr, int64_t nelem, int64_t num_col, const char* parameters, const DatasetHandle reference,
                              DatasetHandle* out);

/* dataset/LightGBMDataset.scala:115,149,163 / :29 — field_name in {label, weight, init_score, group} */
int LGBM_DatasetSetField(DatasetHandle handle, const char* field_name, const void* field_data, int num_element, int type);
int LGBM_DatasetGetField(DatasetHandle handle, const char* field_name, int* out_len, const void** out_ptr, int* out_type);

/* dataset/LightGBMDataset.scala:54 / :65 / :182 / :191 */
int LGBM_DatasetGetNumData(DatasetHandle handle, int* out);
int LGBM_DatasetGetNumFeature(DatasetHandle handle, int* out);
int LGBM_DatasetSetFeatureNames(DatasetHandle handle, const char** feature_names, int num_feature_names);
int LGBM_DatasetFree(DatasetHandle handle);

/* ---------------------------------------------------------------- Booster */

/* booster/LightGBMBooster.scala:241 (from TrainUtils.scala:16-29) */
int LGBM_BoosterCreate(const DatasetHandle train_data, const char* parameters, BoosterHandle* out);
/* booster/LightGBMBooster.scala:355-365 (from TrainUtils.scala:98-135) — one boosting iteration (kernels K3-K8) */
int LGBM_BoosterUpdateOneIter(BoosterHandle handle, int* is_finished);
/* booster/LightGBMBooster.scala:372-392 (from TrainUtils.scala:80-86, FObjTrait): one iteration grown from the caller's gradients;
   grad / hess are host arrays of num_data floats; no automatic initial score (upstream GBDT::TrainOneIter with gradients) */
int LGBM_BoosterUpdateOneIterCustom(BoosterHandle handle, const float* grad, const float* hess, int* is_finished);
/* booster/LightGBMBooster.scala:316-323 (learning-rate schedule through the delegate) */
int LGBM_BoosterResetParameter(BoosterHandle handle, const char* parameters);
/* booster/LightGBMBooster.scala:272-278 (SWIG wrapper LGBM_BoosterSaveModelToStringSWIG grows the buffer and retries) */
int LGBM_BoosterSaveModelToString(BoosterHandle handle, int start_iteration, int num_iteration, int feature_importance_type,
                                  int64_t buffer_len, int64_t* out_len, char* out_str);
/* booster/LightGBMBooster.scala:480-499 (dumpModel; the SWIG wrapper LGBM_BoosterDumpModelSWIG grows the buffer): the model as
   JSON in upstream's GBDT::DumpModel layout; same buffer protocol as SaveModelToString.  Needs no GPU. */
int LGBM_BoosterDumpModel(BoosterHandle handle, int start_iteration, int num_iteration, int feature_importance_type,
                          int64_t buffer_len, int64_t* out_len, char* out_str);
/* booster/LightGBMBooster.scala:163,193,173,183 */
int LGBM_BoosterGetNumClasses(BoosterHandle handle, int* out_len);
int LGBM_BoosterGetNumFeature(BoosterHandle handle, int* out_len);
int LGBM_BoosterNumModelPerIteration(BoosterHandle handle, int* out_tree_per_iteration);
int LGBM_BoosterNumberOfTotalModel(BoosterHandle handle, int* out_models);
int LGBM_BoosterGetCurrentIteration(BoosterHandle handle, int* out_iteration);
/* booster/LightGBMBooster.scala:331-350 — raw training scores (data_idx 0) */
int LGBM_BoosterGetPredict(BoosterHandle handle, int data_idx, int64_t* out_len, double* out_result);
/* booster/LightGBMBooster.scala:155 */
int LGBM_BoosterFree(BoosterHandle handle);

/* ---------------------------------------------------------------- Network */

/* NetworkManager.scala:195-218 — machines = "host:port,host:port,...".  num_machines == 1 is a no-op.  Otherwise the
 * list is used once to hand rank 0's NCCL unique id to the other ranks; collectives then run over NVLink. */
int LGBM_NetworkInit(const char* machines, int local_listen_port, int listen_time_out, int num_machines);
/* BasePartitionTask.scala:282 */
int LGBM_NetworkFree(void);

/* ---------------------------------------------------------------- Extensions (no lightgbmlib counterpart)
 * Introspection used by the parity tests and by bench.py; not called by the Scala layer. */
int LGBM_B200_DatasetGetFeatureInfo(DatasetHandle handle, int feature, int* num_bin, int* missing_type, int* default_bin,
                                    int* most_freq_bin, int* is_trivial, double* min_val, double* max_val);
int LGBM_B200_DatasetGetBinUpperBounds(DatasetHandle handle, int feature, double* out);
int LGBM_B200_DatasetGetNumUsedFeature(DatasetHandle handle, int* out);      /* non-trivial features */
int LGBM_B200_DatasetGetNumColumns(DatasetHandle handle, int* out);         /* uint8 columns of the bin matrix (a feature with > 256 bins has several) */
int LGBM_B200_DatasetCopyBinColumn(DatasetHandle handle, int inner_feature, uint8_t* out);
int LGBM_B200_DatasetCopyBinColumn16(DatasetHandle handle, int inner_feature, uint16_t* out);   /* bins of the k-th non-trivial feature */
int LGBM_B200_DatasetCopyRowMajor(DatasetHandle handle, int64_t row, int64_t nrows, int* out_stride, uint8_t* out);
int LGBM_B200_BoosterGetGradients(BoosterHandle handle, float* grad, float* hess);
int LGBM_B200_BoosterGetQuantGradients(BoosterHandle handle, int8_t* grad, int8_t* hess, double* scales);
int LGBM_B200_BoosterRootHistogram(BoosterHandle handle, int64_t* out /* F_used*256*2 */);
int LGBM_B200_BoosterTimeRootHistogram(BoosterHandle handle, int repeats, float* out_ms);
int LGBM_B200_BoosterSetHistTiming(BoosterHandle handle, int enable);
/* measurement hook for the row-binning kernel (K2): copies `nrow` raw rows (the dataset's own first rows, row-major f64) to the
   device once and times the production kernel re-binning them `repeats` times from that resident copy; ms per pass */
int LGBM_B200_DatasetTimeBinning(DatasetHandle handle, const double* rows, int32_t nrow, int repeats, float* out_ms);
/* per-kernel-type CUDA-event timing of the training kernels (measurement only: it serialises the launches) and the booster's
   set-up timings, returned as one JSON object (same buffer convention as LGBM_BoosterSaveModelToString) */
int LGBM_B200_BoosterSetProfile(BoosterHandle handle, int enable);
int LGBM_B200_BoosterGetProfile(BoosterHandle handle, int64_t buffer_len, int64_t* out_len, char* out_str);
/* all-gather of one value per rank over the TCP star LGBM_NetworkInit opens (CPU-only; for tests of the multi-process path) */
int LGBM_B200_NetworkAllGatherInt64(int64_t mine, int64_t* out, int* out_rank, int* out_world);
int LGBM_B200_BoosterGetStats(BoosterHandle handle, int64_t* hist_rows_visited, double* hist_ms, int64_t* kernel_launches,
                              int64_t* hist_launches);
/* batched best-first growth statistics summed over all trees: dependent rounds, leaves expanded, splits kept */
int LGBM_B200_BoosterGetGrowthStats(BoosterHandle handle, int64_t* rounds, int64_t* expansions, int64_t* splits,
                                    int64_t* hist_kernel_launches);
/* the TCP hand-off LGBM_NetworkInit uses for the NCCL unique id (CPU-only; for tests of the multi-process path) */
int LGBM_B200_NetworkBootstrapBlob(const char* machines, int local_listen_port, int listen_time_out, uint8_t* blob, int32_t bytes,
                                   int* out_rank);
/* CUDA events on the booster's own stream (the stream every kernel of the training loop is launched on) */
int LGBM_B200_BoosterTimerStart(BoosterHandle handle);
int LGBM_B200_BoosterTimerStop(BoosterHandle handle, float* out_ms);
int LGBM_B200_BoosterPredictRaw(BoosterHandle handle, const double* data, int32_t nrow, int32_t ncol, double* out);

/* booster/LightGBMBooster.scala:254-258 (TrainUtils.scala:22-24, continued training from modelString): the trees of `other_handle`
   go in front of this booster's (upstream GBDT::MergeFrom); training scores are not changed by the merge */
int LGBM_BoosterMerge(BoosterHandle handle, BoosterHandle other_handle);
/* ---- validation data and metrics: what early stopping evaluates every iteration
 * (LGBM/TrainUtils.scala:137-169 through LGBM/booster/LightGBMBooster.scala:283-314; the validation dataset is created from the
 * training dataset's bin mappers, LGBM/StreamingPartitionTask.scala:154-171).  AddValidData must precede the first iteration.
 * data_idx 0 = training data, i >= 1 = i-th validation set.  Metrics: binary_logloss, binary_error, auc, l2, rmse, ndcg@k
 * (`metric=` list, default by objective; `eval_at` for ndcg).  Device reductions with a fixed order; values are local to the
 * process. */
int LGBM_BoosterAddValidData(BoosterHandle handle, const DatasetHandle valid_data);
int LGBM_BoosterGetEvalCounts(BoosterHandle handle, int* out_len);
int LGBM_BoosterGetEvalNames(BoosterHandle handle, const int len, int* out_len, const size_t buffer_len, size_t* out_buffer_len,
                             char** out_strs);
int LGBM_BoosterGetEval(BoosterHandle handle, int data_idx, int* out_len, double* out_results);

/* ---- scoring.  Prediction runs on the GPU of the calling process (one thread per row walks the trees);
 * a booster loaded from a model string can only predict and save.
 * LoadModelFromString: LGBM/booster/LightGBMBooster.scala:41-49 (the model string is shipped to the executors and a booster is
 *   rebuilt there).  Supported: what this library writes (num_class = 1, numerical splits, no linear trees).
 * PredictForMat / PredictForMatSingleRow: the C functions behind the SWIG overload LGBM_BoosterPredictForMatSingle
 *   (LGBM/booster/LightGBMBooster.scala:520-557, called per row from the scoring UDF with thread-local output buffers).
 *   predict_type 0 = normal (sigmoid for binary), 1 = raw score, 2 = leaf index; 3 (feature contributions) returns -1.
 *   The single-row form costs one kernel launch and one device round trip per call: batch rows through PredictForMat. */
int LGBM_BoosterLoadModelFromString(const char* model_str, int* out_num_iterations, BoosterHandle* out);
/* LGBM/booster/LightGBMBooster.scala:501-508 (getFeatureImportances): importance_type 0 = split counts, 1 = total gains;
 * out_results has one entry per feature (LGBM_BoosterGetNumFeature). */
int LGBM_BoosterFeatureImportance(BoosterHandle handle, int num_iteration, int importance_type, double* out_results);
int LGBM_BoosterCalcNumPredict(BoosterHandle handle, int num_row, int predict_type, int start_iteration, int num_iteration,
                               int64_t* out_len);
int LGBM_BoosterPredictForMat(BoosterHandle handle, const void* data, int data_type, int32_t nrow, int32_t ncol, int is_row_major,
                              int predict_type, int start_iteration, int num_iteration, const char* parameter, int64_t* out_len,
                              double* out_result);
/* sparse rows (LGBM/booster/LightGBMBooster.scala:520-538): expanded to dense on the host (absent = 0.0), then the same kernel */
int LGBM_BoosterPredictForCSR(BoosterHandle handle, const void* indptr, int indptr_type, const int32_t* indices, const void* data,
                              int data_type, int64_t nindptr, int64_t nelem, int64_t num_col, int predict_type, int start_iteration,
                              int num_iteration, const char* parameter, int64_t* out_len, double* out_result);
int LGBM_BoosterPredictForCSRSingleRow(BoosterHandle handle, const void* indptr, int indptr_type, const int32_t* indices,
                                       const void* data, int data_type, int64_t nindptr, int64_t nelem, int64_t num_col,
                                       int predict_type, int start_iteration, int num_iteration, const char* parameter,
                                       int64_t* out_len, double* out_result);
int LGBM_BoosterPredictForMatSingleRow(BoosterHandle handle, const void* data, int data_type, int ncol, int is_row_major,
                                       int predict_type, int start_iteration, int num_iteration, const char* parameter,
                                       int64_t* out_len, double* out_result);

/* No counterpart in the reference.  Histogram (kernel K4, root-leaf case) for callers that may not allocate device memory:
 * an executor plugin living inside an external memory pool or a fixed per-task GPU budget, or a loop that must not take the
 * driver's allocation lock.  Its scratch (int64 accumulation area + descriptors, about 1 MB) is a static array in the
 * library's device image: held by every GPU that loads the library until exit, shared by all calls (calls overlapping on one
 * GPU race; this call clears what must start at zero), capacity num_features <= 256 (larger is refused with an error).  All
 * pointers are device pointers; stream is a cudaStream_t or NULL.  The first call on a device resolves the array's address —
 * make it before any stream capture.  Calls no cudaMalloc / cudaFree. */
int LGBM_B200_RootHistogramNoAlloc(const uint8_t* d_rows, int32_t row_stride, const uint16_t* d_gh, int32_t num_rows,
                                   int32_t num_features, int32_t num_grad_quant_bins, int64_t* d_out_hist, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* LIGHTGBM_B200_C_API_H_ */
